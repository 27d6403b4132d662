// bpcg.cu -- Bramble-Pasciak CG, device-resident engine for both reference variants:
//   solvers/bramblepasciak_new.py:24-253 (mode 2) and bramble_pasciak_cg.py:65-148 (mode 1; same iterates,
//   SURVEY.md 8(c)(iii)).  The iteration is v2's (one A apply per step, A*s0 and A*w0 by recurrence):
//   B1  SpMV B^T s1, row epilogue:  s0 = beta s0 + w0 (:240-241 of the previous pass) ; As0 = beta As0 + z0 ;
//                                   tmp0 = As0 + B^T s1 ; tmp1 = k dA tmp0 ; tmp4 = tmp1 - s0   (:205-209,:212)
//   B2  SpMV A tmp1 -> tmp2,        fused  sum s0 (tmp2 - tmp0)                        (:210,:218,:222)
//   B3  SpMV B tmp4 -> tmp3,        fused  sum s1 tmp3            -> alpha             (:213,:219,:222,:226)
//   B4  u += a s ; d -= a v ; w0 -= a tmp1 ; w1 -= a dM tmp3 ; z0 -= a tmp2,  fused <w,d> -> beta, err, exit (:228-247)
//   B5  s1 = beta s1 + w1  (pressure part of :240-241; n_p entries)
// z_old of :215 is folded away: beta*As0 + z_old - alpha*tmp2 == beta*As0 + z0(after :238).
// Traffic beyond the three SpMVs: 8*(24 n_u + 13 n_p) bytes per iteration.
// Only diagonal preA / preM are fused here (the reference's BDDC / AMG live in NGSolve: BASELINE.md 4).
#include <math.h>

#include "cond.cuh"
#include "p2p.cuh"

namespace nsb {

struct BpcgState {
    double k, wd, wdn, alpha, beta, as_part, err0, tol_abs;
    double red[3];
    int it, maxsteps, stop, mode, status, nupd;
};

__device__ void bpcg_after_as(BpcgState* S, double part_u, double part_p) {
    const double as_s = part_u + part_p;   // :222
    S->alpha = S->wd / as_s;               // :226
    if (!isfinite(S->alpha)) {
        S->stop = 1;
        S->status = NSB_BREAKDOWN;
    }
}

__device__ void bpcg_after_wd(BpcgState* S, double wdn, double* hist) {
    const int it = S->it;
    const double wd = S->wd;
    S->wdn = wdn;                          // :235
    S->beta = wdn / wd;                    // :236
    S->nupd = it + 1;
    const double err_next = sqrt(fabs(wdn));
    hist[it + 1] = err_next;
    if (!isfinite(wdn)) {
        S->stop = 1;
        S->status = NSB_BREAKDOWN;
        return;
    }
    if (S->mode == 2) {
        const double err = sqrt(fabs(wd));                   // :243  (value from before this update)
        if (err < S->tol_abs) { S->stop = 1; S->status = NSB_CONVERGED; }          // :246
        else if (it + 1 >= S->maxsteps) { S->stop = 1; S->status = NSB_MAXSTEPS; } // :248
    } else {
        if (it + 1 >= S->maxsteps) { S->stop = 1; S->status = NSB_MAXSTEPS; }      // bramble_pasciak_cg.py:110,144
        else if (err_next < S->tol_abs) { S->stop = 1; S->status = NSB_CONVERGED; } // bramble_pasciak_cg.py:119
    }
    S->it = it + 1;
    S->wd = wdn;                           // :221
}

struct BpcgPost {   // scalar step run inside the peer-memory all-reduce kernel
    BpcgState* S;
    int which;
    double* hist;
    __device__ void operator()() const {
        if (S->stop) return;
        if (which == 0) bpcg_after_as(S, S->red[0], S->red[1]);
        else bpcg_after_wd(S, S->red[2], hist);
    }
};

__global__ void k_bpcg_scalar(BpcgState* S, int which, double* hist) {
    if (S->stop) return;
    if (which == 0) bpcg_after_as(S, S->red[0], S->red[1]);
    else bpcg_after_wd(S, S->red[2], hist);
}

struct BpcgB1Op {
    BpcgState* S;
    double* __restrict__ As0;
    const double* __restrict__ z0;
    const double* __restrict__ dA;
    double* __restrict__ s0;
    const double* __restrict__ w0;
    double* __restrict__ tmp0;
    double* __restrict__ tmp1;
    double* __restrict__ tmp4;
    PushOut push;   // row-partitioned: the interface rows of tmp1 / tmp4 go straight into the neighbours' landing buffers (areas 0 / 1)
    __device__ const PushOut& push_out() const { return push; }
    __device__ bool skip() const { return S->stop != 0; }
    struct Pre {
        double as0, z, d, s, w;
    };
    __device__ Pre load(int64_t i) const { return Pre{As0[i], z0[i], dA[i], s0[i], w0[i]}; }
    __device__ void apply(int64_t i, double ax, const Pre& p, double (&)[1]) const {
        const double as0 = S->beta * p.as0 + p.z;
        const double t0 = as0 + ax;
        const double t1 = S->k * (p.d * t0);
        const double sn = S->beta * p.s + p.w;       // :240-241 of the previous pass, velocity part (s0 = w0 before
        s0[i] = sn;                                  // the first pass and beta = 0 then, so this is the identity)
        As0[i] = as0;
        tmp0[i] = t0;
        tmp1[i] = t1;
        tmp4[i] = t1 - sn;
    }
    __device__ void row(int64_t i, double ax, double (&acc)[1]) const { apply(i, ax, load(i), acc); }
    __device__ void row_pushing(int64_t i, double ax, double (&acc)[1]) const {   // rows of a chunk that holds interface rows
        apply(i, ax, load(i), acc);
        push_store<true>(push, i, tmp1[i], tmp4[i]);
    }
    __device__ void finish(const double (&)[1]) const {}
};

// condensed variant of B1: the preconditioner is a chain of SpMVs (cond_apply_pre), so tmp1 / tmp4 are not formed here
struct BpcgB1cOp {
    BpcgState* S;
    double* __restrict__ As0;
    const double* __restrict__ z0;
    double* __restrict__ s0;
    const double* __restrict__ w0;
    double* __restrict__ tmp0;
    __device__ bool skip() const { return S->stop != 0; }
    __device__ void row(int64_t i, double ax, double (&)[1]) const {
        s0[i] = S->beta * s0[i] + w0[i];             // :240-241 of the previous pass
        const double as0 = S->beta * As0[i] + z0[i]; // :205
        As0[i] = as0;
        tmp0[i] = as0 + ax;                          // :207
    }
    __device__ void finish(const double (&)[1]) const {}
};

struct BpcgB2Op {
    BpcgState* S;
    double* __restrict__ tmp2;
    const double* __restrict__ tmp0;
    const double* __restrict__ s0;
    int red_kind;   // kRedLocal: the total is kept for B3's scalar step; else: it goes to the all-reduce landing zone
    __device__ bool skip() const { return S->stop != 0; }
    struct Pre {
        double s, t0;
    };
    __device__ Pre load(int64_t i) const { return Pre{s0[i], tmp0[i]}; }
    __device__ void apply(int64_t i, double ax, const Pre& p, double (&acc)[1]) const {
        tmp2[i] = ax;
        acc[0] = fma(p.s, ax - p.t0, acc[0]);
    }
    __device__ void row(int64_t i, double ax, double (&acc)[1]) const { apply(i, ax, load(i), acc); }
    __device__ void finish(const double (&tot)[1]) const {
        if (red_kind == kRedLocal) S->as_part = tot[0];
        else S->red[0] = tot[0];
    }
};

struct BpcgB3Op {
    BpcgState* S;
    double* __restrict__ tmp3;
    const double* __restrict__ s1;
    RedMode red;
    __device__ bool skip() const { return S->stop != 0; }
    struct Pre {
        double s;
    };
    __device__ Pre load(int64_t j) const { return Pre{s1[j]}; }
    __device__ void apply(int64_t j, double ax, const Pre& p, double (&acc)[1]) const {
        tmp3[j] = ax;
        acc[0] = fma(p.s, ax, acc[0]);
    }
    __device__ void row(int64_t j, double ax, double (&acc)[1]) const { apply(j, ax, load(j), acc); }
    // last block, all threads: <s,v> = red[0] (A SpMV, earlier kernel) + red[1] summed over the ranks, then alpha (:222-226)
    using BlockFinish = void;
    __device__ void finish_block(const double (&tot)[1]) const {
        if (red.kind == kRedLocal) {
            if (threadIdx.x == 0) bpcg_after_as(S, S->as_part, tot[0]);
            return;
        }
        if (threadIdx.x == 0) S->red[1] = tot[0];
        BpcgState* St = S;
        reduce_finish(red, &S->red[0], 2, [St]() { bpcg_after_as(St, St->red[0], St->red[1]); });
    }
    __device__ void finish(const double (&tot)[1]) const { S->red[1] = tot[0]; }
};

struct BpcgVecs {
    double *u0, *u1, *d0, *d1, *w0, *w1, *s0, *s1, *z0;
    const double *tmp0, *tmp1, *tmp2, *tmp3, *dM;
    int64_t n_u, n_p;
};

__global__ void __launch_bounds__(kBlock) k_bpcg_b4(BpcgState* S, BpcgVecs V, double* hist, double* partials,
                                                    unsigned int* ticket, RedMode red) {
    pdl_enter();
    if (S->stop) return;
    const double a = S->alpha;
    double acc[1] = {0.0};
    const int64_t n = V.n_u + V.n_p;
    for (int64_t i = (int64_t)blockIdx.x * blockDim.x + threadIdx.x; i < n; i += (int64_t)gridDim.x * blockDim.x) {
        if (i < V.n_u) {
            const double t2 = V.tmp2[i];
            V.u0[i] += a * V.s0[i];                         // :228
            const double d = V.d0[i] + (-a) * (t2 - V.tmp0[i]);   // :229 with v0 = tmp2 - tmp0 (:218)
            const double w = V.w0[i] + (-a) * V.tmp1[i];    // :232
            V.d0[i] = d;
            V.w0[i] = w;
            V.z0[i] -= a * t2;                              // :238
            acc[0] = fma(w, d, acc[0]);
        } else {
            const int64_t j = i - V.n_u;
            const double t3 = V.tmp3[j];
            V.u1[j] += a * V.s1[j];
            const double d = V.d1[j] + (-a) * t3;           // :229 with v1 = tmp3 (:219)
            const double w = V.w1[j] + (-a) * (V.dM ? V.dM[j] * t3 : t3);   // :233
            V.d1[j] = d;
            V.w1[j] = w;
            acc[0] = fma(w, d, acc[0]);
        }
    }
    if (grid_sum<1>(acc, partials, ticket)) {   // last block: <w,d> over the ranks, then beta, err, exit test (:235-247)
        if (threadIdx.x == 0) S->red[2] = acc[0];
        reduce_finish(red, &S->red[2], 1, [S, hist]() { bpcg_after_wd(S, S->red[2], hist); });
    }
}

// s1 = beta s1 + w1 (:240-241, pressure part; the velocity part rides in B1's epilogue of the next pass)
// Row-partitioned: the interface rows of the new s1 go straight into the neighbours' landing buffers (the B^T SpMV of the next
// pass takes them from there).
__global__ void __launch_bounds__(kBlock) k_bpcg_b5(const BpcgState* S, BpcgVecs V, PushOut P) {
    pdl_enter();
    if (S->stop) return;
    const double b = S->beta;
    u64 e = 0;
    if (P.nruns > 0) {
        e = push_begin(P);
        push_wait_consumed(P, e);
    }
    for (int64_t j = (int64_t)blockIdx.x * blockDim.x + threadIdx.x; j < V.n_p; j += (int64_t)gridDim.x * blockDim.x) {
        const double v = b * V.s1[j] + V.w1[j];
        V.s1[j] = v;
        if (P.nruns > 0 && push_tile(P, j)) push_store<false>(P, j, v, 0.0);
    }
    if (P.nruns > 0) push_end(P, e);
}

__global__ void k_bpcg_init_scalar(BpcgState* S, double k, double wdn_u, const double* wdn_dev, int maxsteps, double tol,
                                   int rel_err, int mode, double* hist) {
    const double wdn = wdn_dev[0] + wdn_dev[1];   // :185
    (void)wdn_u;
    S->k = k;
    S->wdn = wdn;
    S->wd = wdn;
    S->alpha = 0.0;
    S->beta = 0.0;                        // first pass: As0 = 0*As0 + z0 = A s0 (:202-203)
    S->as_part = 0.0;
    S->err0 = sqrt(fabs(wdn));            // :187
    S->tol_abs = tol * (rel_err ? S->err0 : 1.0);   // :246
    S->it = 0;
    S->maxsteps = maxsteps;
    S->mode = mode;
    S->nupd = 0;
    S->stop = 0;
    S->status = NSB_MAXSTEPS;
    hist[0] = S->err0;
    if (wdn == 0.0) { S->stop = 1; S->status = NSB_ZERO_RESIDUAL; }           // :191-192
    else if (!isfinite(wdn)) { S->stop = 1; S->status = NSB_BREAKDOWN; }
    else if (maxsteps <= 0) { S->stop = 1; S->status = NSB_MAXSTEPS; }
    else if (mode == 1 && S->err0 < S->tol_abs) { S->stop = 1; S->status = NSB_CONVERGED; }
}

// y = M x with the halo exchange the operator needs
static int apply(Context& c, nsb_mat_s* m, double* x, double* y) {
    NSB_CHECK(halo_exchange(c, m->halo, x));
    return spmv_plain(c, m, x, y);
}

}  // namespace nsb

using namespace nsb;

static int bpcg_core(nsb_ctx ctx, nsb_mat A, nsb_cond_s* K, nsb_mat B, nsb_mat BT, nsb_pre preA, nsb_pre preM, double k_scale,
                     const double* f, const double* g, double* sol, int maxsteps, double tol, int initialize,
                     int rel_err, int mode, double* history, int* updates, double* seconds, int* status) {
    NSB_REQUIRE(ctx && A && B && BT && preA && f && g && sol && history, "nsb_bpcg: NULL argument");
    NSB_REQUIRE(mode == 1 || mode == 2, "nsb_bpcg: mode must be 1 or 2");
    NSB_REQUIRE(isfinite(k_scale) && k_scale > 0.0, "nsb_bpcg: k_scale must be positive and finite");
    Context& c = ctx->c;
    NSB_CUDA(cudaSetDevice(c.device));
    NvtxRange nvtx_call("nsb_bpcg");
    const int64_t n_u = A->rows, n_p = B->rows;
    const int64_t nxu = A->cols, nxp = BT->cols;
    NSB_REQUIRE(B->cols == nxu, "nsb_bpcg: B has %lld columns, A has %lld", (long long)B->cols, (long long)nxu);
    NSB_REQUIRE(BT->rows == n_u, "nsb_bpcg: BT must have n_u rows");
    NSB_REQUIRE(preA->n == n_u, "nsb_bpcg: preA size mismatch");
    if (preM) NSB_REQUIRE(preM->n == n_p, "nsb_bpcg: preM size mismatch");
    NSB_REQUIRE(maxsteps >= 0, "nsb_bpcg: maxsteps < 0");
    // preA: diagonal (fused into the SpMV epilogues) or a block smoother (MypreA's Jacobi / Gauss-Seidel part,
    // NavierStokesSIMPLE_iterative.py:364-383), applied as its own kernels between B1 and B2
    const bool general_pre = preA->kind != 0;
    const double* dA = general_pre ? nullptr : preA->d.p;
    if (preM) NSB_REQUIRE(preM->kind == 0, "nsb_bpcg: preM must be diagonal");
    const double* dM = preM ? preM->d.p : nullptr;
    if (K) NSB_REQUIRE(mode == 2, "nsb_bpcg_condensed: v2 semantics only");
    if (K) NSB_REQUIRE(!general_pre, "nsb_bpcg_condensed: diagonal preA_unscaled only");
    if (K) NSB_REQUIRE(K->n == n_u && K->nx == nxu, "nsb_bpcg_condensed: operator and B disagree on the velocity vector (%lld | %lld entries "
                       "against %lld | %lld): row-partitioned, all velocity-column operators share one halo plan",
                       (long long)K->n, (long long)K->nx, (long long)n_u, (long long)nxu);
    DevBuf<double> cr, cy;   // scratch of the condensed preconditioner chain (SpMV inputs: owned + halo entries)
    if (K) {
        NSB_CHECK(cr.allocs(c, nxu));
        NSB_CHECK(cy.allocs(c, nxu));
    }
    // matA x -> y and preA x -> y (k * preA_unscaled, extended harmonically when condensed: :8-21,:84-103,:122)
    auto applyA = [&](double* x, double* y) -> int {
        if (K) return cond_apply_A<0>(c, K, x, PlainRowOp{y});
        return apply(c, A, x, y);
    };
    auto applyPre = [&](double* x, double* y) -> int {   // condensed and row-partitioned: x's halo tail is filled (x holds nxu entries)
        if (K) return cond_apply_pre(c, K, x, dA, k_scale, cr.p, cy.p, y);
        if (general_pre) {
            NSB_CHECK(pre_apply_device(c, preA, x, y));
            return vec_lincomb(c, y, k_scale, y, 0, nullptr, 0, nullptr, n_u);
        }
        return vec_diagmul(c, y, k_scale, dA, x, n_u);
    };

    DevBuf<double> u0, u1, d0, d1, w0, w1, s0, s1, z0, As0, tmp0, tmp1, tmp2, tmp3, tmp4, df, dg, fnew, gnew, hist;
    DevBuf<BpcgState> st;
    for (DevBuf<double>* b : {&u0, &s0, &tmp0, &tmp1, &tmp4, &df}) NSB_CHECK(b->allocs(c, nxu));   // (df: SpMV input of the condensed chain)
    for (DevBuf<double>* b : {&d0, &w0, &z0, &As0, &tmp2, &fnew}) NSB_CHECK(b->allocs(c, n_u));
    for (DevBuf<double>* b : {&u1, &s1}) NSB_CHECK(b->allocs(c, nxp));
    for (DevBuf<double>* b : {&d1, &w1, &tmp3, &dg, &gnew}) NSB_CHECK(b->allocs(c, n_p > 0 ? n_p : 1));
    NSB_CHECK(hist.allocs(c, maxsteps + 2));
    NSB_CHECK(st.allocs(c, 1));
    BpcgState* S = st.p;

    // ---- set-up, solvers/bramblepasciak_new.py:124-189 (not condensed)
    NSB_CHECK(copy_in(c, df.p, f, n_u));
    NSB_CHECK(copy_in(c, dg.p, g, n_p));
    NSB_CHECK(applyPre(df.p, tmp0.p));                                            // :122,:128-129
    NSB_CHECK(applyA(tmp0.p, fnew.p));
    NSB_CHECK(vec_lincomb(c, fnew.p, 1.0, fnew.p, -1.0, df.p, 0, nullptr, n_u));  // :130
    NSB_CHECK(apply(c, B, tmp0.p, gnew.p));
    NSB_CHECK(vec_lincomb(c, gnew.p, 1.0, gnew.p, -1.0, dg.p, 0, nullptr, n_p));  // :133  (== tmp5, :177)
    NSB_CHECK(vec_lincomb(c, w0.p, 1.0, tmp0.p, 0, nullptr, 0, nullptr, n_u));    // pr0 (:173) parked in w0
    if (initialize) {                                                              // :138-139
        NSB_CHECK(vec_set(c, u0.p, 0.0, nxu));
        NSB_CHECK(vec_set(c, u1.p, 0.0, nxp));
    } else {
        NSB_CHECK(copy_in(c, u0.p, sol, n_u));
        NSB_CHECK(copy_in(c, u1.p, sol + n_u, n_p));
    }
    NSB_CHECK(applyA(u0.p, tmp0.p));                                               // :160
    NSB_CHECK(apply(c, BT, u1.p, tmp2.p));
    NSB_CHECK(vec_lincomb(c, tmp0.p, 1.0, tmp0.p, 1.0, tmp2.p, 0, nullptr, n_u));
    NSB_CHECK(applyPre(tmp0.p, tmp1.p));                                           // :162
    NSB_CHECK(applyA(tmp1.p, tmp2.p));                                             // :163
    NSB_CHECK(vec_lincomb(c, tmp4.p, 1.0, tmp1.p, -1.0, u0.p, 0, nullptr, n_u));   // :165
    NSB_CHECK(apply(c, B, tmp4.p, tmp3.p));                                        // :166
    NSB_CHECK(vec_lincomb(c, d0.p, 1.0, fnew.p, -1.0, tmp2.p, 1.0, tmp0.p, n_u)); // :169
    NSB_CHECK(vec_lincomb(c, d1.p, 1.0, gnew.p, -1.0, tmp3.p, 0, nullptr, n_p));  // :170
    NSB_CHECK(vec_lincomb(c, w0.p, 1.0, w0.p, -1.0, tmp1.p, 0, nullptr, n_u));    // :182
    NSB_CHECK(vec_lincomb(c, gnew.p, 1.0, gnew.p, -1.0, tmp3.p, 0, nullptr, n_p)); // tmp5 - tmp3
    NSB_CHECK(vec_diagmul(c, w1.p, 1.0, dM, gnew.p, n_p));                         // :179,:183 (preM linear)
    NSB_CHECK(vec_dot(c, w0.p, d0.p, n_u, &S->red[0]));                            // :185
    NSB_CHECK(vec_dot(c, w1.p, d1.p, n_p, &S->red[1]));
    k_bpcg_init_scalar<<<1, 1, 0, c.stream>>>(S, k_scale, 0.0, S->red, maxsteps, tol, rel_err, mode, hist.p);
    c.launches++;
    NSB_CHECK(vec_lincomb(c, s0.p, 1.0, w0.p, 0, nullptr, 0, nullptr, n_u));      // :189
    NSB_CHECK(vec_lincomb(c, s1.p, 1.0, w1.p, 0, nullptr, 0, nullptr, n_p));
    NSB_CHECK(applyA(s0.p, As0.p));                                                // :202
    NSB_CHECK(vec_lincomb(c, z0.p, 1.0, As0.p, 0, nullptr, 0, nullptr, n_u));     // :203

    EventPair ev;
    NSB_CHECK(ev.create());
    NSB_CUDA(cudaEventRecord(ev.e0, c.stream));                                    // timer_its (:195-196)

    BpcgVecs V{u0.p, u1.p, d0.p, d1.p, w0.p, w1.p, s0.p, s1.p, z0.p, tmp0.p, tmp1.p, tmp2.p, tmp3.p, dM, n_u, n_p};
    const int vg = vec_grid(n_u + n_p, c.num_sms);
    // row-partitioned with the peer-memory path: tmp1 and tmp4 are pushed together and the A and B SpMVs each wait
    // for their landing area inside the kernel (interior chunks first), no wait/copy kernels in between
    const bool fused_AB = A->halo && A->halo == B->halo && halo_fused_ok(c, A) && halo_fused_ok(c, B) && !(K || general_pre);
    // ... and the vectors are pushed by the kernels that PRODUCE them: tmp1 / tmp4 from the row epilogue of the B^T SpMV, the new
    // s1 from the pressure-direction kernel; both reductions end in the last block of the kernel that formed the partial sums
    // (peer all-reduce + scalar step).  A row-partitioned iteration is then the same five launches as on one GPU.
    const RedMode red = red_mode(c);
    const int rk = red.kind;
    PushOut push_u, push_p;
    memset(&push_u, 0, sizeof(push_u));
    memset(&push_p, 0, sizeof(push_p));
    if (fused_AB && push_fused_ok(c, A) && push_fused_ok(c, B) && BT->panels.size() == 1) push_u = p2p_push_out(A->halo);
    if (!(K || general_pre) && BT->halo && push_fused_ok(c, BT)) push_p = p2p_push_out(BT->halo);
    if (push_p.nruns > 0) NSB_CHECK(p2p_halo_push(c, BT->halo, s1.p));   // the first pass finds s1's interface entries pushed
    auto reduce_unfused = [&](double* vals, int count, int which) -> int {   // NCCL fallback: all-reduce + scalar kernel
        if (rk != kRedStore) return NSB_OK;
        NSB_CHECK(allreduce_sum(c, vals, count));
        k_bpcg_scalar<<<1, 1, 0, c.stream>>>(S, which, hist.p);
        c.launches++;
        return NSB_OK;
    };
    auto launch_iteration = [&]() -> int {
        if (K || general_pre) {
            NSB_CHECK(halo_exchange(c, BT->halo, s1.p));
            NSB_CHECK(spmv_launch<0>(c, BT, s1.p, BpcgB1cOp{S, As0.p, z0.p, s0.p, w0.p, tmp0.p}));
            NSB_CHECK(applyPre(tmp0.p, tmp1.p));                                                   // :209
            NSB_CHECK(vec_lincomb(c, tmp4.p, 1.0, tmp1.p, -1.0, s0.p, 0, nullptr, n_u));           // :212
            if (K) NSB_CHECK((cond_apply_A<1>(c, K, tmp1.p, BpcgB2Op{S, tmp2.p, tmp0.p, s0.p, rk})));   // :210,:218,:222
            else {
                NSB_CHECK(halo_exchange(c, A->halo, tmp1.p));   // (no-op on one GPU)
                NSB_CHECK(spmv_launch<1>(c, A, tmp1.p, BpcgB2Op{S, tmp2.p, tmp0.p, s0.p, rk}));
            }
        } else {
            const BpcgB1Op b1{S, As0.p, z0.p, dA, s0.p, w0.p, tmp0.p, tmp1.p, tmp4.p, push_u};
            if (push_p.nruns > 0) NSB_CHECK(spmv_launch_halo_in<0>(c, BT, s1.p, b1, 0));      // s1 was pushed by b5 (or before the loop)
            else NSB_CHECK(spmv_exchange_launch<0>(c, BT, s1.p, b1));
            // tmp1 and tmp4 travel in the velocity halo, in one exchange when A and B share the halo plan
            if (fused_AB) {
                if (push_u.nruns == 0) NSB_CHECK(p2p_halo_push(c, A->halo, tmp1.p, tmp4.p));
                NSB_CHECK(spmv_launch_halo_in<1>(c, A, tmp1.p, BpcgB2Op{S, tmp2.p, tmp0.p, s0.p, rk}, 0));
            } else {
                if (A->halo == B->halo) NSB_CHECK(halo_exchange2(c, A->halo, tmp1.p, tmp4.p));
                else NSB_CHECK(halo_exchange(c, A->halo, tmp1.p));
                NSB_CHECK(spmv_launch<1>(c, A, tmp1.p, BpcgB2Op{S, tmp2.p, tmp0.p, s0.p, rk}));
                if (A->halo != B->halo) NSB_CHECK(halo_exchange(c, B->halo, tmp4.p));
            }
        }
        if (K || general_pre) NSB_CHECK(halo_exchange(c, B->halo, tmp4.p));
        if (fused_AB) NSB_CHECK(spmv_launch_halo_in<1>(c, B, tmp4.p, BpcgB3Op{S, tmp3.p, s1.p, red}, 1));
        else NSB_CHECK(spmv_launch<1>(c, B, tmp4.p, BpcgB3Op{S, tmp3.p, s1.p, red}));
        NSB_CHECK(reduce_unfused(&S->red[0], 2, 0));
        NSB_CUDA(launch_k(c, 2, k_bpcg_b4, vg, kBlock, 0, S, V, hist.p, c.partials, c.ticket, red));
        c.launches++;
        NSB_CHECK(reduce_unfused(&S->red[2], 1, 1));
        NSB_CUDA(launch_k(c, 2, k_bpcg_b5, vec_grid(n_p > 0 ? n_p : 1, c.num_sms), kBlock, 0, S, V, push_p));
        c.launches++;
        NSB_CUDA(cudaGetLastError());
        return NSB_OK;
    };

    const int chunk = (n_u + n_p) < 200000 ? 16 : 4;   // iterations in flight between host checks
    NvtxRange nvtx_its("nsb_bpcg: iterations");   // timer_its (:195-196)
    GraphExecHolder gh;
    int64_t graph_launches = 0;
    int rc = NSB_OK;
    if (c.use_graphs && maxsteps > 0)
        rc = capture_batch(c, chunk, [&](int) { return launch_iteration(); }, &gh.exec, &graph_launches, "nsb_bpcg");
    cudaGraphExec_t gexec = gh.exec;
    if (rc == NSB_OK)
        rc = drive_batches(c, &S->stop, maxsteps, chunk, [&]() -> int {
            if (gexec) {
                cudaError_t ge = cudaGraphLaunch(gexec, c.stream);
                if (ge != cudaSuccess) {
                    set_error("nsb_bpcg: graph launch failed: %s", cudaGetErrorString(ge));
                    return NSB_ERR_CUDA;
                }
                c.launches += graph_launches;
                return NSB_OK;
            }
            for (int j = 0; j < chunk; ++j) NSB_CHECK(launch_iteration());
            return NSB_OK;
        });
    NSB_CHECK(rc);
    NSB_CHECK(check_fault(c, "nsb_bpcg"));

    NSB_CUDA(cudaEventRecord(ev.e1, c.stream));
    BpcgState hs;
    NSB_CUDA(cudaMemcpyAsync(&hs, S, sizeof(BpcgState), cudaMemcpyDeviceToHost, c.stream));
    NSB_CUDA(cudaStreamSynchronize(c.stream));
    float ms = 0.f;
    cudaEventElapsedTime(&ms, ev.e0, ev.e1);
    NSB_CHECK(copy_out(c, sol, u0.p, n_u));
    NSB_CHECK(copy_out(c, sol + n_u, u1.p, n_p));
    NSB_CHECK(copy_out(c, history, hist.p, hs.nupd + 1));
    NSB_CUDA(cudaStreamSynchronize(c.stream));
    if (updates) *updates = hs.nupd;
    if (seconds) *seconds = (double)ms * 1e-3;
    if (status) *status = hs.status;
    return NSB_OK;
}

extern "C" int nsb_bpcg(nsb_ctx ctx, nsb_mat A, nsb_mat B, nsb_mat BT, nsb_pre preA, nsb_pre preM, double k_scale,
                        const double* f, const double* g, double* sol, int maxsteps, double tol, int initialize,
                        int rel_err, int mode, double* history, int* updates, double* seconds, int* status) {
    return bpcg_core(ctx, A, nullptr, B, BT, preA, preM, k_scale, f, g, sol, maxsteps, tol, initialize, rel_err, mode, history,
                     updates, seconds, status);
}

extern "C" int nsb_cond_create(nsb_ctx ctx, nsb_mat A, nsb_mat H, nsb_mat HT, nsb_mat inner_solve, nsb_mat inner_matrix,
                               nsb_cond* out) {
    NSB_REQUIRE(ctx && A && H && HT && inner_solve && inner_matrix && out, "nsb_cond_create: NULL argument");
    const int64_t n = A->rows;
    // One GPU: five square operators.  Row-partitioned: this rank's rows of each, columns numbered [owned | halo] over ONE velocity
    // halo plan (an operator that references owned columns only may come without a plan and with n columns).
    nsb_halo_s* plan = nullptr;
    for (nsb_mat m : {A, H, HT, inner_solve, inner_matrix})
        if (m->halo) {
            NSB_REQUIRE(plan == nullptr || plan == m->halo, "nsb_cond_create: the row-partitioned operators must share one halo plan");
            plan = m->halo;
        }
    const int64_t nx = plan ? plan->n_owned + plan->n_halo : n;
    if (plan) NSB_REQUIRE(plan->n_owned == n, "nsb_cond_create: the halo plan owns %lld entries, the operators have %lld rows",
                          (long long)plan->n_owned, (long long)n);
    for (nsb_mat m : {A, H, HT, inner_solve, inner_matrix})
        NSB_REQUIRE(m->rows == n && m->cols == (m->halo ? nx : n),
                    "nsb_cond_create: all five operators must have the same rows and be square (owned + halo columns when row-partitioned)");
    nsb_cond_s* K = new nsb_cond_s();
    K->ctx = &ctx->c;
    K->A = A;
    K->H = H;
    K->HT = HT;
    K->inner_solve = inner_solve;
    K->inner_matrix = inner_matrix;
    K->n = n;
    K->nx = nx;
    int r = K->t.alloc(nx);
    if (r == NSB_OK) r = K->q.alloc(nx);
    if (r != NSB_OK) {
        delete K;
        return r;
    }
    *out = K;
    return NSB_OK;
}

extern "C" int nsb_cond_destroy(nsb_cond k) {
    delete k;
    return NSB_OK;
}

extern "C" int nsb_cond_apply(nsb_ctx ctx, nsb_cond k, const double* x, double* y) {
    NSB_REQUIRE(ctx && k && x && y, "nsb_cond_apply: NULL argument");
    Context& c = ctx->c;
    NSB_CUDA(cudaSetDevice(c.device));
    DevBuf<double> dx, dy;
    NSB_CHECK(dx.allocs(c, k->nx));
    NSB_CHECK(dy.allocs(c, k->n));
    NSB_CHECK(copy_in(c, dx.p, x, k->n));
    NSB_CHECK(cond_apply_A<0>(c, k, dx.p, PlainRowOp{dy.p}));
    NSB_CHECK(copy_out(c, y, dy.p, k->n));
    NSB_CUDA(cudaStreamSynchronize(c.stream));
    return NSB_OK;
}

extern "C" int nsb_bpcg_condensed(nsb_ctx ctx, nsb_cond cond, nsb_mat B, nsb_mat BT, nsb_pre preA, nsb_pre preM, double k_scale,
                                  const double* f, const double* g, double* sol, int maxsteps, double tol, int initialize,
                                  int rel_err, double* history, int* updates, double* seconds, int* status) {
    NSB_REQUIRE(cond, "nsb_bpcg_condensed: NULL operator");
    return bpcg_core(ctx, cond->A, cond, B, BT, preA, preM, k_scale, f, g, sol, maxsteps, tol, initialize, rel_err, 2, history,
                     updates, seconds, status);
}
