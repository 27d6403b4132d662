// cg.cu -- preconditioned CG, CG-Lanczos eigenvalue estimate and modified Gram-Schmidt.
//   nsb_cg            NGSolve CGSolver as used at templates/NavierStokesSIMPLE_iterative.py:93,128 and the
//                     Jacobi-CG that stands in for heat.mat.Inverse (heat.py:72)
//   nsb_lanczos_eigs  EigenValues_Preconditioner (bramble_pasciak_cg.py:70; solvers/bramblepasciak_new.py:115)
//   nsb_orthonormalize  orthonormalization.py:5-16
// CG iteration = 3 kernels:  C1 SpMV q = A s, fused <s,q> -> alpha ;  C2 u += a s, d -= a q, w = C d,
// fused <w,d> -> beta, err, exit ;  C3 s = beta s + w.
#include <math.h>

#include <algorithm>

#include "cg.cuh"
#include "cond.cuh"
#include "p2p.cuh"

namespace nsb {

__device__ void cg_after_as(CgState* S, double as_s, double* alphas) {
    S->alpha = S->wd / as_s;
    if (alphas) alphas[S->it] = S->alpha;
    if (!isfinite(S->alpha)) { S->stop = 1; S->status = NSB_BREAKDOWN; }
}

__device__ void cg_after_wd(CgState* S, double wdn, double* hist, double* betas) {
    const int it = S->it;
    S->wdn = wdn;
    S->beta = wdn / S->wd;
    if (betas) betas[it] = S->beta;
    const double err = sqrt(fabs(wdn));
    if (hist) hist[it + 1] = err;
    S->it = it + 1;
    S->wd = wdn;
    if (!isfinite(wdn) || (betas && wdn <= 0.0)) { S->stop = 1; S->status = NSB_BREAKDOWN; return; }
    if (!betas && err <= S->tol_abs) { S->stop = 1; S->status = NSB_CONVERGED; }
    else if (it + 1 >= S->maxsteps) { S->stop = 1; S->status = NSB_MAXSTEPS; }
}

struct CgPost {   // scalar step run inside the peer-memory all-reduce kernel
    CgState* S;
    int which;
    double *hist, *alphas, *betas;
    __device__ void operator()() const {
        if (S->stop) return;
        if (which == 0) cg_after_as(S, S->red[0], alphas);
        else cg_after_wd(S, S->red[1], hist, betas);
    }
};

__global__ void k_cg_scalar(CgState* S, int which, double* hist, double* alphas, double* betas) {
    if (S->stop) return;
    if (which == 0) cg_after_as(S, S->red[0], alphas);
    else cg_after_wd(S, S->red[1], hist, betas);
}

struct CgC1Op {
    CgState* S;
    double* __restrict__ q;
    const double* __restrict__ s;
    double* alphas;
    RedMode red;
    __device__ bool skip() const { return S->stop != 0; }
    struct Pre {
        double s;
    };
    __device__ Pre load(int64_t i) const { return Pre{s[i]}; }
    __device__ void apply(int64_t i, double ax, const Pre& p, double (&acc)[1]) const {
        q[i] = ax;
        acc[0] = fma(p.s, ax, acc[0]);
    }
    __device__ void row(int64_t i, double ax, double (&acc)[1]) const { apply(i, ax, load(i), acc); }
    using BlockFinish = void;   // last block, all threads: <s, A s> over the ranks, then alpha
    __device__ void finish_block(const double (&tot)[1]) const {
        if (threadIdx.x == 0) S->red[0] = tot[0];
        CgState* St = S;
        double* al = alphas;
        reduce_finish(red, &S->red[0], 1, [St, al]() { cg_after_as(St, St->red[0], al); });
    }
    __device__ void finish(const double (&tot)[1]) const { S->red[0] = tot[0]; }
};

__global__ void __launch_bounds__(kBlock) k_cg_c2(CgState* S, double* __restrict__ u, const double* __restrict__ s,
                                                  double* __restrict__ d, const double* __restrict__ q,
                                                  const double* __restrict__ dinv, double* __restrict__ w, int64_t n,
                                                  double* hist, double* betas, double* partials, unsigned int* ticket,
                                                  RedMode red, bool defer_pre) {
    pdl_enter();
    if (S->stop) return;
    const double a = S->alpha;
    double acc[1] = {0.0};
    for (int64_t i = (int64_t)blockIdx.x * blockDim.x + threadIdx.x; i < n; i += (int64_t)gridDim.x * blockDim.x) {
        if (u) u[i] += a * s[i];
        const double di = d[i] + (-a) * q[i];
        d[i] = di;
        if (defer_pre) continue;   // a general preconditioner forms w = C d and <w,d> with its own kernels behind this one
        const double wi = dinv ? dinv[i] * di : di;
        w[i] = wi;
        acc[0] = fma(wi, di, acc[0]);
    }
    if (defer_pre) return;
    if (grid_sum<1>(acc, partials, ticket)) {   // last block: <w,d> over the ranks, then beta, err and the stop test
        if (threadIdx.x == 0) S->red[1] = acc[0];
        reduce_finish(red, &S->red[1], 1, [S, hist, betas]() { cg_after_wd(S, S->red[1], hist, betas); });
    }
}

// Row-partitioned: the interface rows of the new s go straight into the neighbours' landing buffers (next pass's A s).
__global__ void __launch_bounds__(kBlock) k_cg_c3(const CgState* S, double* __restrict__ s, const double* __restrict__ w,
                                                  int64_t n, PushOut P) {
    pdl_enter();
    if (S->stop) return;
    const double b = S->beta;
    u64 e = 0;
    if (P.nruns > 0) {
        e = push_begin(P);
        push_wait_consumed(P, e);
    }
    for (int64_t i = (int64_t)blockIdx.x * blockDim.x + threadIdx.x; i < n; i += (int64_t)gridDim.x * blockDim.x) {
        const double v = b * s[i] + w[i];
        s[i] = v;
        if (P.nruns > 0 && push_tile(P, i)) push_store<false>(P, i, v, 0.0);
    }
    if (P.nruns > 0) push_end(P, e);
}

__global__ void k_cg_init_scalar(CgState* S, int maxsteps, double tol, double* hist, int lanczos) {
    const double wdn = S->red[0];
    S->wd = S->wdn = wdn;
    S->alpha = S->beta = 0.0;
    S->err0 = sqrt(fabs(wdn));
    S->tol_abs = tol * S->err0;
    S->it = 0;
    S->maxsteps = maxsteps;
    S->stop = 0;
    S->status = NSB_MAXSTEPS;
    S->record = lanczos;
    if (hist) hist[0] = S->err0;
    if (S->err0 == 0.0) { S->stop = 1; S->status = NSB_CONVERGED; }
    else if (!isfinite(wdn)) { S->stop = 1; S->status = NSB_BREAKDOWN; }
    else if (maxsteps <= 0) { S->stop = 1; S->status = NSB_MAXSTEPS; }
}

__global__ void __launch_bounds__(kBlock) k_lanczos_start(double* __restrict__ r, int64_t n) {
    for (int64_t i = (int64_t)blockIdx.x * blockDim.x + threadIdx.x; i < n; i += (int64_t)gridDim.x * blockDim.x)
        r[i] = 1.0 + 0.5 * sin((double)(i + 1));
}

struct ResidualOp2 {  // d = rhs - A u
    const double* __restrict__ rhs;
    double* __restrict__ out;
    __device__ bool skip() const { return false; }
    __device__ void row(int64_t r, double ax, double (&)[1]) const { out[r] = rhs[r] - ax; }
    __device__ void finish(const double (&)[1]) const {}
};

static int cg_run(Context& c, nsb_mat_s* A, nsb_pre_s* pre, CgBuffers& b, bool want_u, int maxsteps, double tol,
                  bool lanczos, int chunk, bool (*host_check)(void*, const CgState&), void* check_arg, CgState* out_state,
                  float* out_ms, nsb_cond_s* K = nullptr, CgGraph* cache = nullptr) {
    const int64_t n = A->rows;
    NvtxRange nvtx_call(lanczos ? "nsb_lanczos_eigs: CG-Lanczos" : "nsb_cg");
    // a block smoother as preconditioner is applied between the update kernel and the <w,d> reduction (unfused)
    const bool general_pre = pre != nullptr && pre->kind != 0;
    const double* dinv = (pre && !general_pre) ? pre->d.p : nullptr;
    const RedMode red = red_mode(c);
    CgState* S = b.st.p;
    const int vg = vec_grid(n, c.num_sms);
    double* u = want_u ? b.u.p : nullptr;
    double* hist = lanczos ? nullptr : b.hist.p;
    double* alphas = lanczos ? b.alphas.p : nullptr;
    double* betas = lanczos ? b.betas.p : nullptr;
    // w = C d ; s = w ; wd = <w, d>
    if (general_pre) NSB_CHECK(pre_apply_device(c, pre, b.d.p, b.w.p));
    else NSB_CHECK(vec_diagmul(c, b.w.p, 1.0, dinv, b.d.p, n));
    NSB_CHECK(vec_lincomb(c, b.s.p, 1.0, b.w.p, 0, nullptr, 0, nullptr, n));
    NSB_CHECK(vec_dot(c, b.w.p, b.d.p, n, &S->red[0]));
    k_cg_init_scalar<<<1, 1, 0, c.stream>>>(S, maxsteps, tol, hist, lanczos ? 1 : 0);
    c.launches++;
    EventPair ev;
    NSB_CHECK(ev.create());
    NSB_CUDA(cudaEventRecord(ev.e0, c.stream));
    // Row-partitioned over peer memory: both reductions end in the last block of the kernel that formed the partial sums and C3
    // ships the interface rows of the s it writes -- three launches per iteration, as on one GPU.
    PushOut push_s;
    memset(&push_s, 0, sizeof(push_s));
    if (!K && A->halo && push_fused_ok(c, A)) push_s = p2p_push_out(A->halo);
    if (push_s.nruns > 0) NSB_CHECK(p2p_halo_push(c, A->halo, b.s.p));   // the first pass finds s's interface entries pushed
    auto reduce_unfused = [&](double* vals, int which) -> int {   // NCCL fallback: all-reduce + scalar kernel
        if (red.kind != kRedStore) return NSB_OK;
        NSB_CHECK(allreduce_sum(c, vals, 1));
        k_cg_scalar<<<1, 1, 0, c.stream>>>(S, which, hist, alphas, betas);
        c.launches++;
        return NSB_OK;
    };
    auto launch_iteration = [&]() -> int {
        const CgC1Op c1{S, b.q.p, b.s.p, alphas, red};
        if (K) NSB_CHECK((cond_apply_A<1>(c, K, b.s.p, c1)));   // condensed: the chain exchanges its own inputs (cond.cuh)
        else if (push_s.nruns > 0) NSB_CHECK(spmv_launch_halo_in<1>(c, A, b.s.p, c1, 0));
        else NSB_CHECK(spmv_exchange_launch<1>(c, A, b.s.p, c1));
        NSB_CHECK(reduce_unfused(&S->red[0], 0));
        NSB_CUDA(launch_k(c, 2, k_cg_c2, vg, kBlock, 0, S, u, b.s.p, b.d.p, b.q.p, dinv, b.w.p, n, hist, betas, c.partials, c.ticket, red,
                          general_pre));
        c.launches++;
        if (general_pre) {   // w = P d, then <w,d>
            NSB_CHECK(pre_apply_device(c, pre, b.d.p, b.w.p));
            NSB_CHECK(vec_dot(c, b.w.p, b.d.p, n, &S->red[1]));
            k_cg_scalar<<<1, 1, 0, c.stream>>>(S, 1, hist, alphas, betas);
            c.launches++;
        } else NSB_CHECK(reduce_unfused(&S->red[1], 1));
        NSB_CUDA(launch_k(c, 2, k_cg_c3, vg, kBlock, 0, S, b.s.p, b.w.p, n, push_s));
        c.launches++;
        NSB_CUDA(cudaGetLastError());
        return NSB_OK;
    };
    // the captured batch of `chunk` iterations; a caller that solves repeatedly on the same buffers (time stepper) keeps it
    CgGraph local_graph;
    CgGraph& G = cache ? *cache : local_graph;
    int rc = NSB_OK;
    if (c.use_graphs && maxsteps > 0 && (!G.exec || G.chunk != chunk)) {
        G.reset();
        const int64_t before = c.launches;
        cudaGraph_t graph = nullptr;
        NSB_CUDA(cudaStreamBeginCapture(c.stream, cudaStreamCaptureModeThreadLocal));
        for (int j = 0; j < chunk && rc == NSB_OK; ++j) rc = launch_iteration();
        cudaError_t ce = cudaStreamEndCapture(c.stream, &graph);
        G.launches = c.launches - before;
        G.chunk = chunk;
        c.launches = before;
        if (rc == NSB_OK && ce != cudaSuccess) {
            set_error("cg: graph capture failed: %s", cudaGetErrorString(ce));
            rc = NSB_ERR_CUDA;
        }
        if (rc == NSB_OK && (ce = cudaGraphInstantiate(&G.exec, graph, 0)) != cudaSuccess) {
            set_error("cg: graph instantiation failed: %s", cudaGetErrorString(ce));
            G.exec = nullptr;
            rc = NSB_ERR_CUDA;
        }
        if (graph) cudaGraphDestroy(graph);
    }
    cudaGraphExec_t gexec = (c.use_graphs && maxsteps > 0) ? G.exec : nullptr;
    const int64_t graph_launches = G.launches;
    CgState hs;
    memset(&hs, 0, sizeof(hs));
    auto batch = [&]() -> int {
        if (gexec) {
            if (cudaGraphLaunch(gexec, c.stream) != cudaSuccess) {
                set_error("cg: graph launch failed");
                return NSB_ERR_CUDA;
            }
            c.launches += graph_launches;
            return NSB_OK;
        }
        for (int j = 0; j < chunk; ++j) NSB_CHECK(launch_iteration());
        return NSB_OK;
    };
    int launched = 0;
    // plain CG: one batch ahead of the stop test (common.cuh).  Lanczos keeps the batch-by-batch loop: its host-side
    // Ritz-value test decides after every batch and the number of steps it takes is part of the result.
    if (rc == NSB_OK && !host_check) {
        rc = drive_batches(c, &S->stop, maxsteps, chunk, batch);
        launched = maxsteps;
    }
    while (rc == NSB_OK && launched < maxsteps) {
        rc = batch();
        launched += chunk;
        if (rc != NSB_OK) break;
        NSB_CUDA(cudaMemcpyAsync(&hs, S, sizeof(CgState), cudaMemcpyDeviceToHost, c.stream));
        NSB_CUDA(cudaStreamSynchronize(c.stream));
        if (hs.stop) break;
        if (host_check && host_check(check_arg, hs)) break;
    }
    NSB_CHECK(rc);
    NSB_CHECK(check_fault(c, "cg"));
    NSB_CUDA(cudaEventRecord(ev.e1, c.stream));
    NSB_CUDA(cudaMemcpyAsync(&hs, S, sizeof(CgState), cudaMemcpyDeviceToHost, c.stream));
    NSB_CUDA(cudaStreamSynchronize(c.stream));
    float ms = 0.f;
    cudaEventElapsedTime(&ms, ev.e0, ev.e1);
    *out_state = hs;
    if (out_ms) *out_ms = ms;
    return NSB_OK;
}

// ---- symmetric tridiagonal extreme eigenvalues by Sturm bisection (host) -------------------------
static int sturm_count(const std::vector<double>& d, const std::vector<double>& e, int m, double x) {
    int cnt = 0;
    double q = d[0] - x;
    if (q < 0) cnt++;
    for (int i = 1; i < m; ++i) {
        const double qq = (q == 0.0) ? 1e-300 : q;
        q = d[i] - x - e[i - 1] * e[i - 1] / qq;
        if (q < 0) cnt++;
    }
    return cnt;  // number of eigenvalues < x
}

static double tridiag_kth(const std::vector<double>& d, const std::vector<double>& e, int m, int k) {
    double lo = d[0], hi = d[0];
    for (int i = 0; i < m; ++i) {
        const double r = (i > 0 ? fabs(e[i - 1]) : 0.0) + (i < m - 1 ? fabs(e[i]) : 0.0);
        lo = std::min(lo, d[i] - r);
        hi = std::max(hi, d[i] + r);
    }
    for (int it = 0; it < 200; ++it) {
        const double mid = 0.5 * (lo + hi);
        if (mid <= lo || mid >= hi) break;
        if (sturm_count(d, e, m, mid) > k) hi = mid;
        else lo = mid;
    }
    return 0.5 * (lo + hi);
}

struct LanczosCheck {
    Context* c;
    CgBuffers* b;
    double tol;
    std::vector<double> alphas, betas, diag, off;
    int used = 0;       // iterations folded into the tridiagonal so far
    double lam_prev = 0.0;
    bool have_prev = false;
    bool converged = false;
    double lam_min = 0.0, lam_max = 0.0;
    int steps = 0;
};

// Fold the newly available CG coefficients into T and apply the oracle's stopping rule iteration by
// iteration (oracle/krylov_oracle.py: lanczos_eigs): stop once lambda_min moved < tol relative, it >= 4.
static bool lanczos_host_check(void* arg, const CgState& hs) {
    LanczosCheck& L = *(LanczosCheck*)arg;
    const int avail = hs.it;
    if (avail <= L.used) return false;
    L.alphas.resize(avail);
    L.betas.resize(avail);
    cudaMemcpy(L.alphas.data() + L.used, L.b->alphas.p + L.used, sizeof(double) * (avail - L.used), cudaMemcpyDeviceToHost);
    cudaMemcpy(L.betas.data() + L.used, L.b->betas.p + L.used, sizeof(double) * (avail - L.used), cudaMemcpyDeviceToHost);
    for (int it = L.used; it < avail; ++it) {
        if (it == 0) L.diag.push_back(1.0 / L.alphas[0]);
        else {
            L.diag.push_back(1.0 / L.alphas[it] + L.betas[it - 1] / L.alphas[it - 1]);
            L.off.push_back(sqrt(L.betas[it - 1]) / L.alphas[it - 1]);
        }
        const int m = (int)L.diag.size();
        const double lam = tridiag_kth(L.diag, L.off, m, 0);
        L.lam_min = lam;
        L.lam_max = tridiag_kth(L.diag, L.off, m, m - 1);
        L.steps = m;
        if (L.have_prev && fabs(lam - L.lam_prev) <= L.tol * fabs(lam) && it >= 4) {
            L.converged = true;
            L.used = avail;
            return true;
        }
        L.lam_prev = lam;
        L.have_prev = true;
    }
    L.used = avail;
    return false;
}

// ---- Gram-Schmidt kernels ------------------------------------------------------------------------
__global__ void __launch_bounds__(kBlock) k_dots2(const double* __restrict__ bi, const double* __restrict__ bj, int64_t n,
                                                  double* out, double* partials, unsigned int* ticket) {
    double acc[2] = {0.0, 0.0};
    for (int64_t i = (int64_t)blockIdx.x * blockDim.x + threadIdx.x; i < n; i += (int64_t)gridDim.x * blockDim.x) {
        const double a = bi[i];
        acc[0] = fma(a, bj[i], acc[0]);
        acc[1] = fma(a, a, acc[1]);
    }
    if (grid_sum<2>(acc, partials, ticket)) {
        if (threadIdx.x == 0) { out[0] = acc[0]; out[1] = acc[1]; }
    }
}

__global__ void __launch_bounds__(kBlock) k_gs_axpy(double* __restrict__ bj, const double* __restrict__ bi,
                                                    const double* __restrict__ dots, int64_t n) {
    const double coef = dots[0] / dots[1];   // orthonormalization.py:11-12
    for (int64_t i = (int64_t)blockIdx.x * blockDim.x + threadIdx.x; i < n; i += (int64_t)gridDim.x * blockDim.x)
        bj[i] -= coef * bi[i];
}

__global__ void __launch_bounds__(kBlock) k_gs_scale(double* __restrict__ bj, const double* __restrict__ nrm2, int64_t n) {
    const double sc = 1.0 / sqrt(nrm2[0]);   // orthonormalization.py:13
    for (int64_t i = (int64_t)blockIdx.x * blockDim.x + threadIdx.x; i < n; i += (int64_t)gridDim.x * blockDim.x)
        bj[i] = sc * bj[i];
}

// ---- column-batched modified Gram-Schmidt --------------------------------------------------------
// The reference orthogonalises column j against b_0 .. b_{j-1} one pair at a time (2 dots + 1 axpy per pair: 5 vector passes).
// Here a column costs three passes whatever j is:
//   dots    d_i = <b_i, b_j> for all i < j from ONE read of the j + 1 vectors; the MODIFIED Gram-Schmidt coefficients follow from
//           them by the recurrence  c_i = d_i - sum_{k<i} c_k G_ki  (G_ki = <b_k, b_i> of the already normalised columns, kept from
//           their own update pass) -- the same numbers the pairwise sweep produces, without re-reading b_j between pairs;
//   update  b_j -= c_0 b_0, -= c_1 b_1, ... in the reference's order, with |b_j|^2 and the cross products <b_i, b_j> of the NEW
//           column accumulated in the same pass;
//   scale   b_j *= 1 / |b_j|.
// Row-partitioned: each reduction is finished by the last block of its kernel (peer all-reduce, p2p.cuh), one per pass instead of
// one per pair.
constexpr int kGsMax = 8;
struct GsPtrs {
    const double* q[kGsMax];
};
struct GsState {
    double G[kGsMax][kGsMax];   // G[k][i], k < i: <b_k, b_i> of this sweep's normalised columns
    double coef[kGsMax];
    double red[kGsMax];
    double sigma;
};

template <int J>
__global__ void __launch_bounds__(kBlock) k_gs_dots(GsPtrs B, const double* __restrict__ bj, int64_t n, GsState* S, double* partials,
                                                    unsigned int* ticket, RedMode red) {
    double acc[J];
#pragma unroll
    for (int k = 0; k < J; ++k) acc[k] = 0.0;
    for (int64_t i = (int64_t)blockIdx.x * blockDim.x + threadIdx.x; i < n; i += (int64_t)gridDim.x * blockDim.x) {
        const double x = bj[i];
#pragma unroll
        for (int k = 0; k < J; ++k) acc[k] = fma(B.q[k][i], x, acc[k]);
    }
    if (grid_sum<J>(acc, partials, ticket)) {
        if (threadIdx.x == 0) {
#pragma unroll
            for (int k = 0; k < J; ++k) S->red[k] = acc[k];
        }
        reduce_finish(red, S->red, J, [S]() {
            for (int i = 0; i < J; ++i) {                      // <b_i, b_j after the projections on b_0 .. b_{i-1}>  /  <b_i, b_i> (= 1)
                double ci = S->red[i];
                for (int k = 0; k < i; ++k) ci -= S->coef[k] * S->G[k][i];
                S->coef[i] = ci;                               // orthonormalization.py:11
            }
        });
    }
}

template <int J>
__global__ void __launch_bounds__(kBlock) k_gs_update(GsPtrs B, double* __restrict__ bj, int64_t n, GsState* S, int j, double* partials,
                                                      unsigned int* ticket, RedMode red) {
    double cf[J > 0 ? J : 1], acc[J + 1];
#pragma unroll
    for (int k = 0; k < J; ++k) cf[k] = S->coef[k];
#pragma unroll
    for (int k = 0; k <= J; ++k) acc[k] = 0.0;
    for (int64_t i = (int64_t)blockIdx.x * blockDim.x + threadIdx.x; i < n; i += (int64_t)gridDim.x * blockDim.x) {
        double x = bj[i];
        double b[J > 0 ? J : 1];
#pragma unroll
        for (int k = 0; k < J; ++k) b[k] = B.q[k][i];
#pragma unroll
        for (int k = 0; k < J; ++k) x -= cf[k] * b[k];         // orthonormalization.py:12, pair after pair
        if (J > 0) bj[i] = x;
        acc[0] = fma(x, x, acc[0]);
#pragma unroll
        for (int k = 0; k < J; ++k) acc[1 + k] = fma(b[k], x, acc[1 + k]);
    }
    if (grid_sum<J + 1>(acc, partials, ticket)) {
        if (threadIdx.x == 0) {
#pragma unroll
            for (int k = 0; k <= J; ++k) S->red[k] = acc[k];
        }
        reduce_finish(red, S->red, J + 1, [S, j]() {
            const double sigma = 1.0 / sqrt(S->red[0]);        // orthonormalization.py:13
            S->sigma = sigma;
            for (int k = 0; k < J; ++k) S->G[k][j] = S->red[1 + k] * sigma;
        });
    }
}

__global__ void __launch_bounds__(kBlock) k_gs_scale_by(double* __restrict__ bj, const GsState* S, int64_t n) {
    const double sc = S->sigma;
    for (int64_t i = (int64_t)blockIdx.x * blockDim.x + threadIdx.x; i < n; i += (int64_t)gridDim.x * blockDim.x) bj[i] = sc * bj[i];
}

template <int J>
static int gs_column(Context& c, const GsPtrs& B, double* bj, int64_t n, GsState* S, int j, int vg, const RedMode& red) {
    if (J > 0) {
        k_gs_dots<(J > 0 ? J : 1)><<<vg, kBlock, 0, c.stream>>>(B, bj, n, S, c.partials, c.ticket, red);
        c.launches++;
    }
    k_gs_update<J><<<vg, kBlock, 0, c.stream>>>(B, bj, n, S, j, c.partials, c.ticket, red);
    k_gs_scale_by<<<vg, kBlock, 0, c.stream>>>(bj, S, n);
    c.launches += 2;
    return NSB_OK;
}

}  // namespace nsb

using namespace nsb;

namespace nsb {
int cg_alloc(Context& c, CgBuffers& b, nsb_mat_s* A, int maxsteps, bool lanczos) {
    const int64_t n = A->rows, nx = A->cols;
    NSB_CHECK(b.u.allocs(c, nx));
    NSB_CHECK(b.s.allocs(c, nx));
    NSB_CHECK(b.d.allocs(c, n));
    NSB_CHECK(b.w.allocs(c, n));
    NSB_CHECK(b.q.allocs(c, n));
    NSB_CHECK(b.hist.allocs(c, maxsteps + 2));
    NSB_CHECK(b.st.allocs(c, 1));
    if (lanczos) {
        NSB_CHECK(b.alphas.allocs(c, maxsteps + 2));
        NSB_CHECK(b.betas.allocs(c, maxsteps + 2));
    }
    return NSB_OK;
}

// CG on device-resident rhs (in b.q) -> b.u; used by nsb_cg and the heat stepper
int cg_solve_device(Context& c, nsb_mat_s* A, nsb_pre_s* pre, CgBuffers& b, bool initialize, int maxsteps, double tol,
                    CgState* hs, float* ms, CgGraph* cache) {
    const int64_t n = A->rows;
    if (initialize) {
        NSB_CHECK(vec_set(c, b.u.p, 0.0, A->cols));
        NSB_CHECK(vec_lincomb(c, b.d.p, 1.0, b.q.p, 0, nullptr, 0, nullptr, n));
    } else {
        NSB_CHECK(halo_exchange(c, A->halo, b.u.p));
        NSB_CHECK(spmv_launch<0>(c, A, b.u.p, ResidualOp2{b.q.p, b.d.p}));
    }
    const int chunk = n < 200000 ? 16 : 4;
    return cg_run(c, A, pre, b, true, maxsteps, tol, false, chunk, nullptr, nullptr, hs, ms, nullptr, cache);
}
}  // namespace nsb

extern "C" int nsb_cg(nsb_ctx ctx, nsb_mat A, nsb_pre pre, const double* rhs, double* sol, int maxsteps, double tol,
                      int initialize, double* history, int* iterations, double* seconds, int* status) {
    NSB_REQUIRE(ctx && A && rhs && sol, "nsb_cg: NULL argument");
    NSB_REQUIRE(maxsteps >= 0, "nsb_cg: maxsteps < 0");
    Context& c = ctx->c;
    NSB_CUDA(cudaSetDevice(c.device));
    const int64_t n = A->rows;
    if (pre) NSB_REQUIRE(pre->n == n, "nsb_cg: preconditioner size mismatch");
    CgBuffers b;
    NSB_CHECK(cg_alloc(c, b, A, maxsteps, false));
    NSB_CHECK(copy_in(c, b.q.p, rhs, n));
    if (!initialize) NSB_CHECK(copy_in(c, b.u.p, sol, n));
    CgState hs;
    float ms = 0.f;
    NSB_CHECK(cg_solve_device(c, A, pre, b, initialize != 0, maxsteps, tol, &hs, &ms));
    NSB_CHECK(copy_out(c, sol, b.u.p, n));
    if (history) NSB_CHECK(copy_out(c, history, b.hist.p, hs.it + 1));
    NSB_CUDA(cudaStreamSynchronize(c.stream));
    if (iterations) *iterations = hs.it;
    if (seconds) *seconds = (double)ms * 1e-3;
    if (status) *status = hs.status;
    return NSB_OK;
}

static int lanczos_core(nsb_ctx ctx, nsb_mat A, nsb_cond_s* K, nsb_pre pre, const double* start, int maxsteps, double tol,
                        double* lam_min, double* lam_max, int* steps) {
    NSB_REQUIRE(ctx && A && lam_min, "nsb_lanczos_eigs: NULL argument");
    NSB_REQUIRE(maxsteps > 0, "nsb_lanczos_eigs: maxsteps must be positive");
    Context& c = ctx->c;
    NSB_CUDA(cudaSetDevice(c.device));
    const int64_t n = A->rows;
    if (pre) NSB_REQUIRE(pre->n == n, "nsb_lanczos_eigs: preconditioner size mismatch");
    CgBuffers b;
    NSB_CHECK(cg_alloc(c, b, A, maxsteps, true));
    if (start) NSB_CHECK(copy_in(c, b.d.p, start, n));
    else {
        k_lanczos_start<<<vec_grid(n, c.num_sms), kBlock, 0, c.stream>>>(b.d.p, n);
        c.launches++;
    }
    LanczosCheck L;
    L.c = &c;
    L.b = &b;
    L.tol = tol;
    CgState hs;
    float ms = 0.f;
    NSB_CHECK(cg_run(c, A, pre, b, false, maxsteps, 0.0, true, 8, lanczos_host_check, &L, &hs, &ms, K));
    if (!L.converged) lanczos_host_check(&L, hs);  // fold whatever the last chunk produced
    NSB_REQUIRE(L.steps > 0, "nsb_lanczos_eigs: Lanczos produced no step (zero start residual?)");
    *lam_min = L.lam_min;
    if (lam_max) *lam_max = L.lam_max;
    if (steps) *steps = L.steps;
    return NSB_OK;
}

// modified Gram-Schmidt on m device-resident vectors of length n (stride n), orthonormalization.py:8-13, pair by pair: the
// reference's pass structure (2 dots + 1 axpy per pair).  Kept for m > 8 and for the NCCL fallback.
static int gs_pairwise(Context& c, double* Q, int64_t n, int m, int tries) {
    DevBuf<double> sc;
    NSB_CHECK(sc.allocs(c, 2));
    const int vg = vec_grid(n, c.num_sms);
    for (int t = 0; t < tries; ++t) {                                   // :8
        for (int j = 0; j < m; ++j) {                                   // :9
            double* bj = Q + (int64_t)j * n;
            for (int i = 0; i < j; ++i) {                               // :10
                const double* bi = Q + (int64_t)i * n;
                k_dots2<<<vg, kBlock, 0, c.stream>>>(bi, bj, n, sc.p, c.partials, c.ticket);
                NSB_CHECK(allreduce_sum(c, sc.p, 2));
                k_gs_axpy<<<vg, kBlock, 0, c.stream>>>(bj, bi, sc.p, n);
                c.launches += 2;
            }
            k_dots2<<<vg, kBlock, 0, c.stream>>>(bj, bj, n, sc.p, c.partials, c.ticket);
            NSB_CHECK(allreduce_sum(c, sc.p, 2));
            k_gs_scale<<<vg, kBlock, 0, c.stream>>>(bj, sc.p, n);
            c.launches += 2;
        }
    }
    return NSB_OK;
}

// orthonormalize(basis, tries): column-batched (three passes per column, see k_gs_dots) for m <= 8
static int gs_device(Context& c, double* Q, int64_t n, int m, int tries) {
    NvtxRange nvtx_call("nsb_orthonormalize");
    static int batched = -1;
    if (batched < 0) {
        const char* e = getenv("NSB_GS_BATCHED");
        batched = (e && atoi(e) == 0) ? 0 : 1;
    }
    const RedMode red = red_mode(c);
    if (!batched || m > kGsMax || red.kind == kRedStore) return gs_pairwise(c, Q, n, m, tries);
    DevBuf<GsState> st;
    NSB_CHECK(st.allocs(c, 1));
    const int vg = vec_grid(n, c.num_sms);
    GsPtrs B;
    for (int k = 0; k < kGsMax; ++k) B.q[k] = Q + (int64_t)(k < m ? k : 0) * n;
    for (int t = 0; t < tries; ++t) {                                   // :8
        for (int j = 0; j < m; ++j) {                                   // :9
            double* bj = Q + (int64_t)j * n;
            int rc = NSB_OK;
            switch (j) {
                case 0: rc = gs_column<0>(c, B, bj, n, st.p, j, vg, red); break;
                case 1: rc = gs_column<1>(c, B, bj, n, st.p, j, vg, red); break;
                case 2: rc = gs_column<2>(c, B, bj, n, st.p, j, vg, red); break;
                case 3: rc = gs_column<3>(c, B, bj, n, st.p, j, vg, red); break;
                case 4: rc = gs_column<4>(c, B, bj, n, st.p, j, vg, red); break;
                case 5: rc = gs_column<5>(c, B, bj, n, st.p, j, vg, red); break;
                case 6: rc = gs_column<6>(c, B, bj, n, st.p, j, vg, red); break;
                default: rc = gs_column<7>(c, B, bj, n, st.p, j, vg, red); break;
            }
            NSB_CHECK(rc);
        }
    }
    NSB_CUDA(cudaGetLastError());
    return check_fault(c, "nsb_orthonormalize");
}

extern "C" int nsb_lanczos_eigs(nsb_ctx ctx, nsb_mat A, nsb_pre pre, const double* start, int maxsteps, double tol,
                                double* lam_min, double* lam_max, int* steps) {
    return lanczos_core(ctx, A, nullptr, pre, start, maxsteps, tol, lam_min, lam_max, steps);
}

extern "C" int nsb_lanczos_eigs_condensed(nsb_ctx ctx, nsb_cond cond, nsb_pre pre, const double* start, int maxsteps,
                                          double tol, double* lam_min, double* lam_max, int* steps) {
    NSB_REQUIRE(cond, "nsb_lanczos_eigs_condensed: NULL operator");
    return lanczos_core(ctx, cond->A, cond, pre, start, maxsteps, tol, lam_min, lam_max, steps);
}

extern "C" int nsb_orthonormalize(nsb_ctx ctx, double* basis, int64_t n, int m, int tries) {
    NSB_REQUIRE(ctx && basis && n >= 0 && m >= 0 && tries >= 0, "nsb_orthonormalize: bad argument");
    Context& c = ctx->c;
    NSB_CUDA(cudaSetDevice(c.device));
    if (m == 0 || n == 0) return NSB_OK;
    const bool on_dev = is_device_ptr(basis);
    DevBuf<double> own;
    double* Q = basis;
    if (!on_dev) {
        NSB_CHECK(own.allocs(c, n * m));
        NSB_CHECK(copy_in(c, own.p, basis, n * m));
        Q = own.p;
    }
    NSB_CHECK(gs_device(c, Q, n, m, tries));
    NSB_CUDA(cudaGetLastError());
    if (!on_dev) NSB_CHECK(copy_out(c, basis, Q, n * m));
    NSB_CUDA(cudaStreamSynchronize(c.stream));
    return NSB_OK;
}

// ---- orthonormalize without any allocation ------------------------------------------------------------------------------
// nsb_orthonormalize_noalloc: the column-batched modified Gram-Schmidt above for callers that must not allocate (code under CUDA
// graph capture, tight serving loops, harnesses that forbid cudaMalloc).  No counterpart in the reference.
// Every byte of scratch the kernels need -- the GsState (Gram entries, coefficients, reduction landing zone),
// the per-block partial sums of the grid reductions and their arrival ticket -- lives in ONE file-scope device array that every
// GPU loading the library holds until exit (kNoAllocScratchBytes = 96 KiB).  In host code the array's name is only a stand-in:
// its device address is taken once per device with cudaGetSymbolAddress (that device current) and passed to the kernels.
// All calls share the array: calls that overlap on the GPU (two streams, two host threads) race on it.
namespace nsb {

constexpr int kNoAllocMaxGrid = 148 * 8;   // blocks of a vector kernel (vec_grid's cap on a 148-SM part)
constexpr size_t kNoAllocStateBytes = (sizeof(GsState) + 255) / 256 * 256;
constexpr size_t kNoAllocPartialBytes = sizeof(double) * (size_t)(kGsMax + 1) * kNoAllocMaxGrid;
constexpr size_t kNoAllocScratchBytes = 96 * 1024;
static_assert(kNoAllocStateBytes + kNoAllocPartialBytes + 256 <= kNoAllocScratchBytes, "no-alloc scratch too small");

static __device__ __align__(256) unsigned char g_noalloc_scratch[kNoAllocScratchBytes];

// device address of g_noalloc_scratch on device `dev` (current), resolved once per device
static int noalloc_scratch_ptr(int dev, unsigned char** out) {
    static unsigned char* cached[64] = {nullptr};
    if (dev < 0 || dev >= 64) {
        set_error("nsb_orthonormalize_noalloc: device ordinal %d out of range", dev);
        return NSB_ERR_INVALID;
    }
    if (!cached[dev]) {
        unsigned char* p = nullptr;
        NSB_CUDA(cudaGetSymbolAddress((void**)&p, g_noalloc_scratch));
        cached[dev] = p;
    }
    *out = cached[dev];
    return NSB_OK;
}

template <int J>
static void gs_column_static(Context& c, const GsPtrs& B, double* bj, int64_t n, GsState* S, double* partials, unsigned int* ticket,
                             int j, int vg, const RedMode& red) {
    if (J > 0) {
        k_gs_dots<(J > 0 ? J : 1)><<<vg, kBlock, 0, c.stream>>>(B, bj, n, S, partials, ticket, red);
        c.launches++;
    }
    k_gs_update<J><<<vg, kBlock, 0, c.stream>>>(B, bj, n, S, j, partials, ticket, red);
    k_gs_scale_by<<<vg, kBlock, 0, c.stream>>>(bj, S, n);
    c.launches += 2;
}

}  // namespace nsb

extern "C" int nsb_orthonormalize_noalloc(nsb_ctx ctx, double* basis, int64_t n, int m, int tries) {
    NSB_REQUIRE(ctx && basis && n >= 0 && m >= 0 && tries >= 0, "nsb_orthonormalize_noalloc: bad argument");
    Context& c = ctx->c;
    NSB_REQUIRE(c.nranks == 1, "nsb_orthonormalize_noalloc: single-GPU contexts only");
    NSB_REQUIRE(m <= NSB_NOALLOC_MAX_VECTORS, "nsb_orthonormalize_noalloc: m = %d vectors exceeds the maximum of %d the static scratch holds",
                m, NSB_NOALLOC_MAX_VECTORS);
    NSB_REQUIRE(n <= NSB_NOALLOC_MAX_LENGTH, "nsb_orthonormalize_noalloc: n = %lld exceeds the maximum length %lld", (long long)n,
                (long long)NSB_NOALLOC_MAX_LENGTH);
    NSB_CUDA(cudaSetDevice(c.device));
    if (m == 0 || n == 0) return NSB_OK;
    NSB_REQUIRE(is_device_ptr(basis), "nsb_orthonormalize_noalloc: basis must be device-resident (a host basis would need a staging allocation)");
    const int vg = vec_grid(n, c.num_sms);
    NSB_REQUIRE(vg <= kNoAllocMaxGrid, "nsb_orthonormalize_noalloc: a grid of %d blocks exceeds the %d the static scratch holds", vg,
                kNoAllocMaxGrid);
    unsigned char* base = nullptr;
    NSB_CHECK(noalloc_scratch_ptr(c.device, &base));
    GsState* S = (GsState*)base;
    double* partials = (double*)(base + kNoAllocStateBytes);
    unsigned int* ticket = (unsigned int*)(base + kNoAllocStateBytes + kNoAllocPartialBytes);
    // the array keeps the last call's contents: the state and the arrival ticket must start at zero
    NSB_CUDA(cudaMemsetAsync(base, 0, kNoAllocStateBytes, c.stream));
    NSB_CUDA(cudaMemsetAsync(ticket, 0, 256, c.stream));
    RedMode red;
    memset(&red, 0, sizeof(red));
    red.kind = kRedLocal;
    GsPtrs B;
    for (int k = 0; k < kGsMax; ++k) B.q[k] = basis + (int64_t)(k < m ? k : 0) * n;
    for (int t = 0; t < tries; ++t) {                                   // orthonormalization.py:8
        for (int j = 0; j < m; ++j) {                                   // :9
            double* bj = basis + (int64_t)j * n;
            switch (j) {
                case 0: gs_column_static<0>(c, B, bj, n, S, partials, ticket, j, vg, red); break;
                case 1: gs_column_static<1>(c, B, bj, n, S, partials, ticket, j, vg, red); break;
                case 2: gs_column_static<2>(c, B, bj, n, S, partials, ticket, j, vg, red); break;
                case 3: gs_column_static<3>(c, B, bj, n, S, partials, ticket, j, vg, red); break;
                case 4: gs_column_static<4>(c, B, bj, n, S, partials, ticket, j, vg, red); break;
                case 5: gs_column_static<5>(c, B, bj, n, S, partials, ticket, j, vg, red); break;
                case 6: gs_column_static<6>(c, B, bj, n, S, partials, ticket, j, vg, red); break;
                default: gs_column_static<7>(c, B, bj, n, S, partials, ticket, j, vg, red); break;
            }
        }
    }
    NSB_CUDA(cudaGetLastError());
    NSB_CUDA(cudaStreamSynchronize(c.stream));
    return NSB_OK;
}

// ---- heat.py time-loop body (exponential-integrator step in a Krylov-like subspace) ----------------
namespace nsb {

constexpr int kHeatMaxDim = 8;

struct BasisPtrs {
    const double* q[kHeatMaxDim];
};

// out[k] = <Q_k, r>, k < m, from one read of r and the m basis vectors (heat.py:111-118: the inner loops
// over `row` are m InnerProducts against the same residual)
template <int M>
__global__ void __launch_bounds__(kBlock) k_multidot(BasisPtrs B, const double* __restrict__ r, int64_t n, double* out,
                                                     double* partials, unsigned int* ticket) {
    double acc[M];
#pragma unroll
    for (int k = 0; k < M; ++k) acc[k] = 0.0;
    for (int64_t i = (int64_t)blockIdx.x * blockDim.x + threadIdx.x; i < n; i += (int64_t)gridDim.x * blockDim.x) {
        const double ri = r[i];
#pragma unroll
        for (int k = 0; k < M; ++k) acc[k] = fma(B.q[k][i], ri, acc[k]);
    }
    if (grid_sum<M>(acc, partials, ticket)) {
        if (threadIdx.x == 0) {
#pragma unroll
            for (int k = 0; k < M; ++k) out[k] = acc[k];
        }
    }
}

// u = sum_k coef[k] Q_k   (heat.py:140-142)
__global__ void __launch_bounds__(kBlock) k_lift(BasisPtrs B, const double* __restrict__ coef, int m, double* __restrict__ u,
                                                 int64_t n) {
    double cf[kHeatMaxDim];
    for (int k = 0; k < m; ++k) cf[k] = coef[k];
    for (int64_t i = (int64_t)blockIdx.x * blockDim.x + threadIdx.x; i < n; i += (int64_t)gridDim.x * blockDim.x) {
        double s = 0.0;
        for (int k = 0; k < m; ++k) s += cf[k] * B.q[k][i];   // same order as the reference's accumulation
        u[i] = s;
    }
}

static int multidot(Context& c, const BasisPtrs& B, int m, const double* r, int64_t n, double* out) {
    const int vg = vec_grid(n, c.num_sms);
    BasisPtrs P = B;
    for (int k = m; k < kHeatMaxDim; ++k) P.q[k] = B.q[0];
    if (m <= 4) k_multidot<4><<<vg, kBlock, 0, c.stream>>>(P, r, n, out, c.partials, c.ticket);
    else k_multidot<8><<<vg, kBlock, 0, c.stream>>>(P, r, n, out, c.partials, c.ticket);
    c.launches++;
    NSB_CUDA(cudaGetLastError());
    return allreduce_sum(c, out, m <= 4 ? 4 : 8);
}

// dense solve A x = b (n x n row-major, partial pivoting), host; A and b are overwritten
static bool dense_solve(std::vector<double>& A, std::vector<double>& b, int n) {
    for (int k = 0; k < n; ++k) {
        int p = k;
        for (int i = k + 1; i < n; ++i)
            if (fabs(A[(size_t)i * n + k]) > fabs(A[(size_t)p * n + k])) p = i;
        if (A[(size_t)p * n + k] == 0.0) return false;
        if (p != k) {
            for (int j = 0; j < n; ++j) std::swap(A[(size_t)k * n + j], A[(size_t)p * n + j]);
            std::swap(b[k], b[p]);
        }
        for (int i = k + 1; i < n; ++i) {
            const double f = A[(size_t)i * n + k] / A[(size_t)k * n + k];
            if (f == 0.0) continue;
            for (int j = k; j < n; ++j) A[(size_t)i * n + j] -= f * A[(size_t)k * n + j];
            b[i] -= f * b[k];
        }
    }
    for (int i = n - 1; i >= 0; --i) {
        double s = b[i];
        for (int j = i + 1; j < n; ++j) s -= A[(size_t)i * n + j] * b[j];
        b[i] = s / A[(size_t)i * n + i];
    }
    return true;
}

}  // namespace nsb

extern "C" int nsb_heat_step(nsb_ctx ctx, nsb_mat K, nsb_mat M, nsb_mat H, nsb_pre preH, double* u, double time_step, int m,
                             int stages, const double* rk_a, const double* rk_b, double cg_tol, int cg_maxsteps,
                             int* cg_iterations, double* Khat_out, double* Mhat_out) {
    NSB_REQUIRE(ctx && K && M && H && u && rk_a && rk_b, "nsb_heat_step: NULL argument");
    NSB_REQUIRE(m >= 1 && m <= kHeatMaxDim, "nsb_heat_step: subspace dimension must be in [1, %d]", kHeatMaxDim);
    NSB_REQUIRE(stages >= 1 && stages <= 64, "nsb_heat_step: stages must be in [1, 64]");
    Context& c = ctx->c;
    NSB_CUDA(cudaSetDevice(c.device));
    NvtxRange nvtx_call("nsb_heat_step");
    const int64_t n = K->rows, nx = K->cols;
    NSB_REQUIRE(M->rows == n && H->rows == n && M->cols == nx && H->cols == nx, "nsb_heat_step: operator sizes differ");
    if (preH) NSB_REQUIRE(preH->n == n, "nsb_heat_step: preconditioner size mismatch");
    const double dt = time_step / m;                                      // heat.py:75
    DevBuf<double> Q, xu, r, sc, coef;
    NSB_CHECK(Q.allocs(c, (int64_t)m * n));
    NSB_CHECK(xu.allocs(c, nx));                                               // temperature with halo room
    NSB_CHECK(r.allocs(c, nx));
    NSB_CHECK(sc.allocs(c, 2 * kHeatMaxDim * kHeatMaxDim + 16));
    NSB_CHECK(coef.allocs(c, kHeatMaxDim));
    CgBuffers b;
    NSB_CHECK(cg_alloc(c, b, H, cg_maxsteps, false));
    NSB_CHECK(copy_in(c, xu.p, u, n));
    NSB_CHECK(vec_lincomb(c, Q.p, 1.0, xu.p, 0, nullptr, 0, nullptr, n));  // :87
    double* d_norm = sc.p + 2 * kHeatMaxDim * kHeatMaxDim;
    NSB_CHECK(vec_dot(c, xu.p, xu.p, n, d_norm));                          // :89
    int total_cg = 0;
    for (int i = 1; i < m; ++i) {                                          // :95-98
        NSB_CHECK(halo_exchange(c, K->halo, xu.p));
        NSB_CHECK(spmv_plain(c, K, xu.p, b.q.p));                          // residual = diffusion * temperature
        CgState hs;
        NSB_CHECK(cg_solve_device(c, H, preH, b, true, cg_maxsteps, cg_tol, &hs, nullptr));   // heat_inverse * residual
        total_cg += hs.it;
        NSB_CHECK(vec_lincomb(c, xu.p, 1.0, xu.p, -dt, b.u.p, 0, nullptr, n));
        NSB_CHECK(vec_lincomb(c, Q.p + (int64_t)i * n, 1.0, xu.p, 0, nullptr, 0, nullptr, n));
    }
    NSB_CHECK(gs_device(c, Q.p, n, m, 3));                                 // :100
    BasisPtrs B;
    for (int k = 0; k < kHeatMaxDim; ++k) B.q[k] = Q.p + (int64_t)(k < m ? k : 0) * n;
    for (int col = 0; col < m; ++col) {                                    // :109-118
        NSB_CHECK(vec_lincomb(c, xu.p, 1.0, Q.p + (int64_t)col * n, 0, nullptr, 0, nullptr, n));
        NSB_CHECK(halo_exchange(c, K->halo, xu.p));
        NSB_CHECK(spmv_plain(c, K, xu.p, r.p));
        NSB_CHECK(multidot(c, B, m, r.p, n, sc.p + (int64_t)col * kHeatMaxDim));
        NSB_CHECK(spmv_plain(c, M, xu.p, r.p));
        NSB_CHECK(multidot(c, B, m, r.p, n, sc.p + (int64_t)(kHeatMaxDim + col) * kHeatMaxDim));
    }
    std::vector<double> h(2 * kHeatMaxDim * kHeatMaxDim + 16);
    NSB_CUDA(cudaMemcpyAsync(h.data(), sc.p, sizeof(double) * h.size(), cudaMemcpyDeviceToHost, c.stream));
    NSB_CUDA(cudaStreamSynchronize(c.stream));
    // host: m x m projected system and the collocation step (heat.py:120-135, runge_kutta_method.py:26-59)
    std::vector<double> Kh((size_t)m * m), Mh((size_t)m * m), evo((size_t)m * m);
    for (int col = 0; col < m; ++col)
        for (int row = 0; row < m; ++row) {
            Kh[(size_t)row * m + col] = h[(size_t)col * kHeatMaxDim + row];
            Mh[(size_t)row * m + col] = h[(size_t)(kHeatMaxDim + col) * kHeatMaxDim + row];
        }
    if (Khat_out) memcpy(Khat_out, Kh.data(), sizeof(double) * (size_t)m * m);
    if (Mhat_out) memcpy(Mhat_out, Mh.data(), sizeof(double) * (size_t)m * m);
    for (int col = 0; col < m; ++col) {                                    // evo = -Mh^-1 Kh, column by column
        std::vector<double> A = Mh, rhs((size_t)m);
        for (int row = 0; row < m; ++row) rhs[row] = -Kh[(size_t)row * m + col];
        NSB_REQUIRE(dense_solve(A, rhs, m), "nsb_heat_step: projected mass matrix is singular");
        for (int row = 0; row < m; ++row) evo[(size_t)row * m + col] = rhs[row];
    }
    const double norm0 = sqrt(h[2 * kHeatMaxDim * kHeatMaxDim]);
    std::vector<double> y((size_t)m, 0.0), My((size_t)m, 0.0);
    y[0] = norm0;                                                          // :131-133
    for (int row = 0; row < m; ++row) My[row] = evo[(size_t)row * m] * norm0;
    const int N = m * stages;
    std::vector<double> coefm((size_t)N * N, 0.0), kk((size_t)N);
    for (int si = 0; si < stages; ++si)
        for (int sj = 0; sj < stages; ++sj)
            for (int row = 0; row < m; ++row)
                for (int col = 0; col < m; ++col)
                    coefm[(size_t)(si * m + row) * N + (sj * m + col)] =
                        ((si == sj && row == col) ? 1.0 : 0.0) - time_step * rk_a[si * stages + sj] * evo[(size_t)row * m + col];
    for (int si = 0; si < stages; ++si)
        for (int row = 0; row < m; ++row) kk[(size_t)si * m + row] = My[row];
    NSB_REQUIRE(dense_solve(coefm, kk, N), "nsb_heat_step: collocation system is singular");
    double nxt[kHeatMaxDim] = {0};
    for (int row = 0; row < m; ++row) {
        double s = 0.0;
        for (int si = 0; si < stages; ++si) s += rk_b[si] * kk[(size_t)si * m + row];
        nxt[row] = y[row] + time_step * s;
    }
    NSB_CUDA(cudaMemcpyAsync(coef.p, nxt, sizeof(double) * kHeatMaxDim, cudaMemcpyHostToDevice, c.stream));
    k_lift<<<vec_grid(n, c.num_sms), kBlock, 0, c.stream>>>(B, coef.p, m, xu.p, n);   // :140-142
    c.launches++;
    NSB_CUDA(cudaGetLastError());
    NSB_CHECK(copy_out(c, u, xu.p, n));
    NSB_CUDA(cudaStreamSynchronize(c.stream));
    if (cg_iterations) *cg_iterations = total_cg;
    return NSB_OK;
}
