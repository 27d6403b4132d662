// minres.cu -- preconditioned MINRES, device-resident.  Drop-in for the loop of minres.py:12-149.
//
// One iteration of the reference is 1 operator apply, 1 preconditioner apply, 2 inner products and
// 9 separate vector sweeps (minres.py:97-118 ~ 26 vector passes).  Here it is three kernels:
//   K1  t = K zh                        fused  sum t_i zh_i                    -> delta   (:97-98)
//   K2  vh' = ig t - delta ig vh - gamma ig_old vh_old ; zh' = C vh'           (:99-101)
//                                       fused  sum zh'_i vh'_i                 -> gamma_new, Givens (:103-113,:119-127)
//   K3  w' = (ig zh - a3 w_old - a2 w)/a1 ; u += c_new eta_old w'              (:115-118)
// The 1/gamma scalings of :73-74,:104-105 are carried as scalars (zh, vh are the unscaled vectors,
// ig = 1/gamma), which removes four sweeps; SpMV + 13 vector passes per iteration.
// All recurrence scalars live in device memory (MinresState) and are produced by the thread that
// finishes each deterministic grid reduction, so the host never has to look at them: the loop is
// launched ahead in chunks (CUDA graph when not distributed) and a stop flag makes the kernels after
// the exit iteration no-ops.  Exit logic reproduces minres.py:96,126,145-146 exactly.
#include <math.h>

#include "p2p.cuh"
#include "spmv.cuh"

namespace nsb {

struct MinresState {
    double gamma, ig, ig_old, delta;
    double c, c_old, s, s_old, eta_old;
    double res, res_old, err0, tol;
    double a1inv, a2, a3, coef_u, igz;  // parameters of K3 for the current iteration
    double red[2];                       // reduction landing zone in distributed mode
    int k, maxsteps, stop, done, status, nhist;
};

// ---- scalar steps (run by one thread: last block of the reduction, or a 1-thread kernel) --------
__device__ void minres_after_delta(MinresState* S, double sum) { S->delta = S->ig * S->ig * sum; }

__device__ void minres_after_gamma(MinresState* S, double sum, double* hist) {
    const double gamma = S->gamma, delta = S->delta;
    const double gamma_new = sqrt(sum);                                   // :103
    const double a0 = S->c * delta - S->c_old * S->s * gamma;             // :107
    const double a1 = sqrt(a0 * a0 + gamma_new * gamma_new);              // :108
    const double a2 = S->s * delta + S->c_old * S->c * gamma;             // :109
    const double a3 = S->s_old * gamma;                                   // :110
    const double c_new = a0 / a1, s_new = gamma_new / a1;                 // :112-113
    S->a1inv = 1.0 / a1;
    S->a2 = a2;
    S->a3 = a3;
    S->coef_u = c_new * S->eta_old;                                       // :118
    S->igz = S->ig;
    const double eta = -s_new * S->eta_old;                               // :119
    const double res = fabs(s_new) * S->res_old;                          // :122
    hist[S->k] = res / S->err0;                                           // :125
    S->nhist = S->k + 1;
    S->res = res;
    if (!(isfinite(res) && isfinite(gamma_new))) {
        S->stop = 1;
        S->status = NSB_BREAKDOWN;
        return;
    }
    if (res < S->tol * S->err0) {                                         // :126
        S->stop = 1;
        S->status = NSB_CONVERGED;
        return;
    }
    S->k += 1;                                                            // :128
    S->eta_old = eta;                                                     // :135-144
    S->s_old = S->s;
    S->s = s_new;
    S->c_old = S->c;
    S->c = c_new;
    S->ig_old = S->ig;
    S->gamma = gamma_new;
    S->ig = 1.0 / gamma_new;
    S->res_old = res;
    if (!(S->k < S->maxsteps + 1)) {                                      // :96
        S->stop = 1;
        S->status = NSB_MAXSTEPS;
    } else if (!(res > S->tol)) {
        S->stop = 1;
        S->status = NSB_ABSTOL;
    }
}

struct MinresPost {   // scalar step run inside the peer-memory all-reduce kernel
    MinresState* S;
    int which;
    double* hist;
    __device__ void operator()() const {
        if (S->stop) return;
        if (which == 0) minres_after_delta(S, S->red[0]);
        else minres_after_gamma(S, S->red[1], hist);
    }
};

__global__ void k_minres_scalar(MinresState* S, int which, double* hist) {
    if (which == 0) {
        if (S->stop) return;
        minres_after_delta(S, S->red[0]);
    } else {
        if (S->stop) return;
        minres_after_gamma(S, S->red[1], hist);
    }
}

// ---- K1 ---------------------------------------------------------------------------------------
struct MinresK1Op {
    MinresState* S;
    double* __restrict__ t;
    const double* __restrict__ zh;
    RedMode red;
    __device__ bool skip() const {
        const bool st = S->stop != 0;
        if (st && blockIdx.x == 0 && threadIdx.x == 0) S->done = 1;
        return st;
    }
    struct Pre {
        double z;
    };
    __device__ Pre load(int64_t r) const { return Pre{zh[r]}; }
    __device__ void apply(int64_t r, double ax, const Pre& p, double (&acc)[1]) const {
        t[r] = ax;
        acc[0] = fma(ax, p.z, acc[0]);
    }
    __device__ void row(int64_t r, double ax, double (&acc)[1]) const { apply(r, ax, load(r), acc); }
    using BlockFinish = void;   // last block, all threads: delta summed over the ranks (:98)
    __device__ void finish_block(const double (&tot)[1]) const {
        if (threadIdx.x == 0) S->red[0] = tot[0];
        MinresState* St = S;
        reduce_finish(red, &S->red[0], 1, [St]() { minres_after_delta(St, St->red[0]); });
    }
    __device__ void finish(const double (&tot)[1]) const { S->red[0] = tot[0]; }
};

// ---- K2 ---------------------------------------------------------------------------------------
__global__ void __launch_bounds__(kBlock) k_minres_k2(MinresState* S, const double* __restrict__ t,
                                                      const double* __restrict__ vh, const double* __restrict__ vh_old,
                                                      const double* __restrict__ dinv, double* __restrict__ vh_new,
                                                      double* __restrict__ zh_new, int64_t n, double* hist,
                                                      double* partials, unsigned int* ticket, RedMode red, bool defer_pre, PushOut P) {
    pdl_enter();
    if (S->stop) return;
    u64 push_e = 0;
    if (P.nruns > 0) {   // row-partitioned: the interface rows of the new zh go straight to the neighbours (next pass's K zh)
        push_e = push_begin(P);
        push_wait_consumed(P, push_e);
    }
    const double ig = S->ig;
    const double c1 = S->delta * ig;            // delta * v      = delta * ig * vh
    const double c2 = S->gamma * S->ig_old;     // gamma * v_old  = gamma * ig_old * vh_old
    double acc[1] = {0.0};
    for (int64_t i = (int64_t)blockIdx.x * blockDim.x + threadIdx.x; i < n; i += (int64_t)gridDim.x * blockDim.x) {
        const double vn = ig * t[i] - c1 * vh[i] - c2 * vh_old[i];
        vh_new[i] = vn;
        if (defer_pre) continue;                // a general preconditioner is applied by its own kernels behind this one
        const double zn = dinv ? dinv[i] * vn : vn;
        zh_new[i] = zn;
        if (P.nruns > 0 && push_tile(P, i)) push_store<false>(P, i, zn, 0.0);
        acc[0] = fma(zn, vn, acc[0]);
    }
    if (defer_pre) return;
    if (grid_sum<1>(acc, partials, ticket)) {   // last block: gamma_new^2 over the ranks, then the Givens step and the stop test
        if (threadIdx.x == 0) S->red[1] = acc[0];
        reduce_finish(red, &S->red[1], 1, [S, hist]() { minres_after_gamma(S, S->red[1], hist); });
    }
    if (P.nruns > 0) push_end(P, push_e);
}

// ---- K3 ---------------------------------------------------------------------------------------
__global__ void __launch_bounds__(kBlock) k_minres_k3(const MinresState* S, const double* __restrict__ zh,
                                                      const double* __restrict__ w_old, const double* __restrict__ w,
                                                      double* __restrict__ w_new, double* __restrict__ u, int64_t n) {
    pdl_enter();
    if (S->done) return;
    const double igz = S->igz, a3 = S->a3, a2 = S->a2, a1inv = S->a1inv, cu = S->coef_u;
    for (int64_t i = (int64_t)blockIdx.x * blockDim.x + threadIdx.x; i < n; i += (int64_t)gridDim.x * blockDim.x) {
        double wn = igz * zh[i] - a3 * w_old[i] - a2 * w[i];   // :115
        wn = a1inv * wn;                                        // :116
        w_new[i] = wn;
        u[i] += cu * wn;                                        // :118
    }
}

// ---- set-up: zh = C vh, sum zh.vh -> gamma and the initial state (:68-92) -------------------------
__global__ void __launch_bounds__(kBlock) k_minres_init(MinresState* S, const double* __restrict__ vh,
                                                        const double* __restrict__ dinv, double* __restrict__ zh,
                                                        int64_t n, double* partials, unsigned int* ticket) {
    double acc[1] = {0.0};
    for (int64_t i = (int64_t)blockIdx.x * blockDim.x + threadIdx.x; i < n; i += (int64_t)gridDim.x * blockDim.x) {
        const double v = vh[i];
        const double z = dinv ? dinv[i] * v : v;
        zh[i] = z;
        acc[0] = fma(z, v, acc[0]);
    }
    if (grid_sum<1>(acc, partials, ticket)) {
        if (threadIdx.x == 0) S->red[0] = acc[0];
    }
}

__global__ void k_minres_init_scalar(MinresState* S, int maxsteps, double tol, double* hist) {
    const double gamma = sqrt(S->red[0]);   // :71
    S->gamma = gamma;
    S->ig = 1.0 / gamma;                    // :73-74 carried as a scalar
    S->ig_old = 0.0;
    S->delta = 0.0;
    S->c = 1.0;
    S->c_old = 1.0;
    S->s = 0.0;
    S->s_old = 0.0;
    S->eta_old = gamma;                     // :83
    S->res = S->res_old = S->err0 = gamma;  // :76-78
    S->tol = tol;
    S->k = 1;
    S->maxsteps = maxsteps;
    S->stop = 0;
    S->done = 0;
    S->status = NSB_MAXSTEPS;
    hist[0] = 1.0;                          // :95
    S->nhist = 1;
    if (!(gamma > 0.0) || !isfinite(gamma)) {   // zero right-hand side divides by zero at :73
        S->stop = 1;
        S->done = 1;
        S->status = NSB_BREAKDOWN;
    } else if (!(1 < maxsteps + 1)) {
        S->stop = 1;
        S->done = 1;
        S->status = NSB_MAXSTEPS;
    } else if (!(gamma > tol)) {
        S->stop = 1;
        S->done = 1;
        S->status = NSB_ABSTOL;
    }
}

struct ResidualRowOp {  // vh = rhs - K u    (:66)
    const double* __restrict__ rhs;
    double* __restrict__ out;
    __device__ bool skip() const { return false; }
    __device__ void row(int64_t r, double ax, double (&)[1]) const { out[r] = rhs[r] - ax; }
    __device__ void finish(const double (&)[1]) const {}
};

}  // namespace nsb

using namespace nsb;

extern "C" int nsb_minres(nsb_ctx ctx, nsb_mat K, nsb_pre pre, const double* rhs, double* sol, int maxsteps, double tol,
                          int initialize, double* history, int* iterations, double* seconds, int* status) {
    NSB_REQUIRE(ctx && K && rhs && sol && history, "nsb_minres: NULL argument");
    NSB_REQUIRE(maxsteps >= 0, "nsb_minres: maxsteps < 0");
    Context& c = ctx->c;
    NSB_CUDA(cudaSetDevice(c.device));
    NvtxRange nvtx_call("nsb_minres");
    const int64_t n = K->rows;                      // owned rows
    const int64_t nx = K->cols;                     // owned + halo
    NSB_REQUIRE(K->halo ? K->halo->n_owned == n : K->cols == K->rows, "nsb_minres: operator must be square");
    if (pre) NSB_REQUIRE(pre->n == n, "nsb_minres: preconditioner size %lld != %lld", (long long)pre->n, (long long)n);
    // pre: diagonal (fused into K2) or any other device preconditioner -- block smoother, auxiliary-space composite (run.py:61-62
    // and stokes_hcurldiv.py:48 configure non-Jacobi ones) -- applied by its own kernels between K2 and the gamma reduction
    const bool general_pre = pre != nullptr && pre->kind != 0;
    const double* dinv = (pre && !general_pre) ? pre->d.p : nullptr;

    DevBuf<double> u, t, v[3], w[3], z[2], hist, drhs;
    DevBuf<MinresState> st;
    NSB_CHECK(u.allocs(c, nx));
    NSB_CHECK(t.allocs(c, n));
    NSB_CHECK(drhs.allocs(c, n));
    for (auto& b : v) NSB_CHECK(b.allocs(c, n));
    for (auto& b : w) NSB_CHECK(b.allocs(c, n));
    for (auto& b : z) NSB_CHECK(b.allocs(c, nx));
    NSB_CHECK(hist.allocs(c, maxsteps + 2));
    NSB_CHECK(st.allocs(c, 1));
    MinresState* S = st.p;
    const int vg = vec_grid(n, c.num_sms);

    // ---- set-up (:50-92)
    NSB_CHECK(copy_in(c, drhs.p, rhs, n));
    if (initialize) {
        NSB_CHECK(vec_set(c, u.p, 0.0, nx));
        NSB_CHECK(vec_lincomb(c, v[1].p, 1.0, drhs.p, 0, nullptr, 0, nullptr, n));
    } else {
        NSB_CHECK(copy_in(c, u.p, sol, n));
        NSB_CHECK(halo_exchange(c, K->halo, u.p));
        NSB_CHECK(spmv_launch<0>(c, K, u.p, ResidualRowOp{drhs.p, v[1].p}));
    }
    NSB_CHECK(vec_set(c, v[0].p, 0.0, n));   // v_old
    NSB_CHECK(vec_set(c, w[0].p, 0.0, n));   // w_old
    NSB_CHECK(vec_set(c, w[1].p, 0.0, n));   // w
    if (general_pre) {
        NSB_CHECK(pre_apply_device(c, pre, v[1].p, z[0].p));                 // z = C v (:68)
        NSB_CHECK(vec_dot(c, z[0].p, v[1].p, n, &S->red[0]));
    } else {
        k_minres_init<<<vg, kBlock, 0, c.stream>>>(S, v[1].p, dinv, z[0].p, n, c.partials, c.ticket);
        c.launches++;
        NSB_CHECK(allreduce_sum(c, S->red, 1));
    }
    k_minres_init_scalar<<<1, 1, 0, c.stream>>>(S, maxsteps, tol, hist.p);
    c.launches++;
    NSB_CUDA(cudaGetLastError());

    EventPair ev;
    NSB_CHECK(ev.create());
    NSB_CUDA(cudaEventRecord(ev.e0, c.stream));

    // Row-partitioned over peer memory: both reductions end in the last block of the kernel that formed the partial sums, and K2
    // ships the interface rows of the zh it writes -- three launches per iteration, as on one GPU.
    const RedMode red = red_mode(c);
    PushOut push_z;
    memset(&push_z, 0, sizeof(push_z));
    if (K->halo && push_fused_ok(c, K) && !general_pre) push_z = p2p_push_out(K->halo);
    if (push_z.nruns > 0) NSB_CHECK(p2p_halo_push(c, K->halo, z[0].p));   // the first pass finds zh's interface entries pushed
    auto reduce_unfused = [&](double* vals, int which) -> int {   // NCCL fallback: all-reduce + scalar kernel
        if (red.kind != kRedStore) return NSB_OK;
        NSB_CHECK(allreduce_sum(c, vals, 1));
        k_minres_scalar<<<1, 1, 0, c.stream>>>(S, which, hist.p);
        c.launches++;
        return NSB_OK;
    };
    // ---- iteration; pointer rotation has period 6 (v,w: 3-cycles, z: 2-cycle)  (:131-133)
    auto launch_iteration = [&](int j) -> int {
        double* v_old = v[j % 3].p;
        double* vc = v[(j + 1) % 3].p;
        double* v_new = v[(j + 2) % 3].p;
        double* w_old = w[j % 3].p;
        double* wc = w[(j + 1) % 3].p;
        double* w_new = w[(j + 2) % 3].p;
        double* zc = z[j % 2].p;
        double* z_new = z[(j + 1) % 2].p;
        if (push_z.nruns > 0) NSB_CHECK(spmv_launch_halo_in<1>(c, K, zc, MinresK1Op{S, t.p, zc, red}, 0));   // zh was pushed by K2
        else NSB_CHECK(spmv_exchange_launch<1>(c, K, zc, MinresK1Op{S, t.p, zc, red}));
        NSB_CHECK(reduce_unfused(&S->red[0], 0));
        NSB_CUDA(launch_k(c, 2, k_minres_k2, vg, kBlock, 0, S, t.p, vc, v_old, dinv, v_new, z_new, n, hist.p, c.partials, c.ticket, red,
                          general_pre, push_z));
        c.launches++;
        if (general_pre) {   // z_new = C v_new (:101), gamma_new^2 = <z_new, v_new> (:103); ignored once the stop flag is up
            NSB_CHECK(pre_apply_device(c, pre, v_new, z_new));
            NSB_CHECK(vec_dot(c, z_new, v_new, n, &S->red[1]));
            k_minres_scalar<<<1, 1, 0, c.stream>>>(S, 1, hist.p);
            c.launches++;
        } else NSB_CHECK(reduce_unfused(&S->red[1], 1));
        NSB_CUDA(launch_k(c, 2, k_minres_k3, vg, kBlock, 0, S, zc, w_old, wc, w_new, u.p, n));
        c.launches++;
        NSB_CUDA(cudaGetLastError());
        return NSB_OK;
    };

    NvtxRange nvtx_its("nsb_minres: iterations");
    const int period = 6;
    const int reps = n < 200000 ? 4 : 1;   // iterations in flight between host checks: period * reps
    GraphExecHolder gh;
    int64_t graph_launches = 0;
    int rc = NSB_OK;
    if (c.use_graphs && maxsteps > 0)
        rc = capture_batch(c, period, [&](int j) { return launch_iteration(j); }, &gh.exec, &graph_launches, "nsb_minres");
    cudaGraphExec_t gexec = gh.exec;
    if (rc == NSB_OK)
        rc = drive_batches(c, &S->stop, maxsteps, period * reps, [&]() -> int {
            for (int r = 0; r < reps; ++r) {
                if (gexec) {
                    cudaError_t ge = cudaGraphLaunch(gexec, c.stream);
                    if (ge != cudaSuccess) {
                        set_error("nsb_minres: graph launch failed: %s", cudaGetErrorString(ge));
                        return NSB_ERR_CUDA;
                    }
                    c.launches += graph_launches;
                } else {
                    for (int j = 0; j < period; ++j) NSB_CHECK(launch_iteration(j));
                }
            }
            return NSB_OK;
        });
    NSB_CHECK(rc);
    NSB_CHECK(check_fault(c, "nsb_minres"));

    NSB_CUDA(cudaEventRecord(ev.e1, c.stream));
    MinresState hs;
    NSB_CUDA(cudaMemcpyAsync(&hs, S, sizeof(MinresState), cudaMemcpyDeviceToHost, c.stream));
    NSB_CUDA(cudaStreamSynchronize(c.stream));
    float ms = 0.f;
    cudaEventElapsedTime(&ms, ev.e0, ev.e1);
    NSB_CHECK(copy_out(c, sol, u.p, n));
    NSB_CHECK(copy_out(c, history, hist.p, hs.nhist));
    NSB_CUDA(cudaStreamSynchronize(c.stream));
    if (iterations) *iterations = hs.nhist - 1;
    if (seconds) *seconds = (double)ms * 1e-3;
    if (status) *status = hs.status;
    return NSB_OK;
}
