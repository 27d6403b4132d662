// simple.cu -- one time step of the SIMPLE-type splitting of templates/NavierStokesSIMPLE_iterative.py, device-resident:
//   DoTimeStep (:427-438)   temp  = conv_operator gfu + f - a gfu
//                           temp2 = invmstar temp            invmstar = (I + H) CGSolver(mstar, pre, 1e-4) (I + HT) + inner_solve  (:93-98)
//                           Project(temp2[0:V.ndof])
//                           gfu  += timestep temp2
//   Project (:440-443)      tmp   = invproj (bproj vel)      invproj  = (I + H) CGSolver(aproj, pre) (I + HT) + inner_solve        (:128-131)
//                           gfup  = tmp[V2.ndof : V2.ndof + Q.ndof]
//                           vel  -= mapV tmp                 mapV = PermutationMatrix(Xproj.ndof, ind): (mapV x)[i] = x[ind[i]]    (:138-144)
// The reference runs this as ~20 NGSolve statements plus two CGSolver calls; here the whole step is one C call on a handle that owns
// its work vectors: every vector statement rides in the row epilogue of the SpMV next to it (temp in the `a` SpMV, (I + HT) and
// (I + H) and "+ inner_solve" in the epilogues of those SpMVs, the CG right-hand sides are written where the solver reads them),
// the gather through mapV and the update of gfu are one kernel, and the captured CG batches are kept across steps.
// Row-partitioned contexts: every operator carries its own halo plan; the exchange before each SpMV and the CG reductions are the
// ones of the Krylov solvers (p2p.cu / dist.cu).  The matrix-free convection operator (:106-113) stays with the caller (NGSolve
// BilinearForm evaluation): it enters as the vector conv_operator * gfu.
#include "cg.cuh"
#include "cond.cuh"

struct nsb_simple_s {
    nsb::Context* ctx = nullptr;
    nsb_mat_s *a = nullptr, *mstar = nullptr, *mH = nullptr, *mHT = nullptr, *m_inner = nullptr;       // not owned
    nsb_mat_s *bproj = nullptr, *aproj = nullptr, *pH = nullptr, *pHT = nullptr, *p_inner = nullptr;   // not owned
    nsb_pre_s *pre_m = nullptr, *pre_p = nullptr;
    int64_t n_x = 0, n_vel = 0, n_proj = 0, p_begin = 0, p_end = 0;
    double dt = 0, prec_m = 1e-4, prec_p = 1e-8;
    int max_m = 200, max_p = 200;
    bool map_needs_halo = false;
    nsb::DevBuf<int32_t> mapV;
    nsb::DevBuf<double> xu, temp, temp2, velx, rp, tmp, conv_d, f_d;
    nsb::CgBuffers cg_m, cg_p;
    nsb::CgGraph g_m, g_p;
};

namespace nsb {

struct SimpleRhsOp {   // temp = conv + f - a u   (:432-434, in the reference's order of accumulation)
    const double* __restrict__ conv;
    const double* __restrict__ f;
    double* __restrict__ out;
    __device__ bool skip() const { return false; }
    __device__ void row(int64_t r, double ax, double (&)[1]) const {
        double t = conv ? conv[r] : 0.0;
        if (f) t += f[r];
        out[r] = t - ax;
    }
    __device__ void finish(const double (&)[1]) const {}
};

// vel -= mapV tmp ; gfu += dt temp2   (:443 and :438 in one pass; entries past n_vel have no projection part)
__global__ void __launch_bounds__(kBlock) k_simple_update(double* __restrict__ gfu, const double* __restrict__ temp2,
                                                          const double* __restrict__ tmp, const int* __restrict__ map, int64_t n_vel,
                                                          int64_t n_x, double dt) {
    for (int64_t i = (int64_t)blockIdx.x * blockDim.x + threadIdx.x; i < n_x; i += (int64_t)gridDim.x * blockDim.x) {
        double v = temp2[i];
        if (i < n_vel) v -= tmp[map[i]];
        gfu[i] += dt * v;
    }
}

// vel -= mapV tmp  (Project called on its own: :415,:426)
__global__ void __launch_bounds__(kBlock) k_simple_correct(double* __restrict__ vel, const double* __restrict__ tmp,
                                                           const int* __restrict__ map, int64_t n_vel) {
    for (int64_t i = (int64_t)blockIdx.x * blockDim.x + threadIdx.x; i < n_vel; i += (int64_t)gridDim.x * blockDim.x)
        vel[i] -= tmp[map[i]];
}

static int64_t cols_or(const nsb_mat_s* m, int64_t dflt) { return m ? m->cols : dflt; }
static int64_t max3(int64_t a, int64_t b, int64_t c) { return a > b ? (a > c ? a : c) : (b > c ? b : c); }

// y = M x with the operator's own halo exchange in front (x has room for M's halo tail)
template <class Op>
static int apply(Context& c, nsb_mat_s* m, double* x, Op op) {
    NSB_CHECK(halo_exchange(c, m->halo, x));
    return spmv_launch<0>(c, m, x, op);
}

// out = (ext @ inv1 @ extT + inner_solve) x with inv1 = CG(M, pre) from zero (:93-98 / :128-131); H == nullptr: inv1 alone.
// x: device vector with room for the halo tails of HT and inner; returns the result pointer in *out (cg.u or `ybuf`).
static int extended_inverse(Context& c, nsb_mat_s* M, nsb_pre_s* pre, nsb_mat_s* H, nsb_mat_s* HT, nsb_mat_s* inner, CgBuffers& cg,
                            CgGraph& graph, double* x, bool x_is_q, double* ybuf, double tol, int maxsteps, double** out, int* its) {
    CgState hs;
    if (!H) {
        if (!x_is_q) NSB_CHECK(vec_lincomb(c, cg.q.p, 1.0, x, 0, nullptr, 0, nullptr, M->rows));
        NSB_CHECK(cg_solve_device(c, M, pre, cg, true, maxsteps, tol, &hs, nullptr, &graph));
        *its = hs.it;
        *out = cg.u.p;
        return NSB_OK;
    }
    NSB_CHECK(apply(c, HT, x, AddInOp{x, cg.q.p}));                       // r = (I + HT) x, written where the solver reads it
    NSB_CHECK(cg_solve_device(c, M, pre, cg, true, maxsteps, tol, &hs, nullptr, &graph));
    *its = hs.it;
    NSB_CHECK(apply(c, H, cg.u.p, AddInOp{cg.u.p, ybuf}));                // y = (I + H) inv1 r
    NSB_CHECK(apply(c, inner, x, AddInOp{ybuf, ybuf}));                   // y += inner_solve x
    *out = ybuf;
    return NSB_OK;
}

// tmp_projection = invproj (bproj vel) (:441); vel: device vector with room for bproj's halo tail.  *tmp_out: the result
// (length n_proj, with the halo tail of aproj filled when mapV reaches into it)
static int project_core(Context& c, nsb_simple_s* s, double* vel, int* its, double** tmp_out) {
    double* t = nullptr;
    if (!s->pH) {
        NSB_CHECK(apply(c, s->bproj, vel, PlainRowOp{s->cg_p.q.p}));
        NSB_CHECK(extended_inverse(c, s->aproj, s->pre_p, nullptr, nullptr, nullptr, s->cg_p, s->g_p, s->cg_p.q.p, true, nullptr, s->prec_p,
                                   s->max_p, &t, its));
    } else {
        NSB_CHECK(apply(c, s->bproj, vel, PlainRowOp{s->rp.p}));
        NSB_CHECK(extended_inverse(c, s->aproj, s->pre_p, s->pH, s->pHT, s->p_inner, s->cg_p, s->g_p, s->rp.p, false, s->tmp.p, s->prec_p,
                                   s->max_p, &t, its));
    }
    if (s->map_needs_halo) NSB_CHECK(halo_exchange(c, s->aproj->halo, t));
    *tmp_out = t;
    return NSB_OK;
}

}  // namespace nsb

using namespace nsb;

extern "C" int nsb_simple_create(nsb_ctx ctx, nsb_mat a, nsb_mat mstar, nsb_pre pre_mstar, nsb_mat mstar_H, nsb_mat mstar_HT,
                                 nsb_mat mstar_inner_solve, nsb_mat bproj, nsb_mat aproj, nsb_pre pre_aproj, nsb_mat aproj_H,
                                 nsb_mat aproj_HT, nsb_mat aproj_inner_solve, const int32_t* map_ind, int64_t n_vel, int64_t p_begin,
                                 int64_t p_end, double timestep, double mstar_precision, int mstar_maxsteps, double proj_precision,
                                 int proj_maxsteps, nsb_simple* out) {
    NSB_REQUIRE(ctx && a && mstar && bproj && aproj && map_ind && out, "nsb_simple_create: NULL argument");
    const int64_t n_x = a->rows, n_proj = aproj->rows;
    NSB_REQUIRE(mstar->rows == n_x, "nsb_simple_create: mstar has %lld rows, a has %lld", (long long)mstar->rows, (long long)n_x);
    NSB_REQUIRE(n_vel >= 0 && n_vel <= n_x, "nsb_simple_create: n_vel out of range");
    NSB_REQUIRE(bproj->rows == n_proj, "nsb_simple_create: bproj must have the rows of aproj");
    NSB_REQUIRE((bproj->halo ? bproj->halo->n_owned : bproj->cols) == n_vel, "nsb_simple_create: bproj must have n_vel (owned) columns");
    NSB_REQUIRE(0 <= p_begin && p_begin <= p_end && p_end <= n_proj, "nsb_simple_create: pressure range outside the projection vector");
    NSB_REQUIRE((mstar_H != nullptr) == (mstar_HT != nullptr) && (mstar_H != nullptr) == (mstar_inner_solve != nullptr),
                "nsb_simple_create: give all of harmonic_extension, harmonic_extension_trans, inner_solve of mstar or none");
    NSB_REQUIRE((aproj_H != nullptr) == (aproj_HT != nullptr) && (aproj_H != nullptr) == (aproj_inner_solve != nullptr),
                "nsb_simple_create: give all of harmonic_extension, harmonic_extension_trans, inner_solve of aproj or none");
    for (nsb_mat m : {mstar_H, mstar_HT, mstar_inner_solve})
        if (m) NSB_REQUIRE(m->rows == n_x, "nsb_simple_create: the condensation operators of mstar must have %lld rows", (long long)n_x);
    for (nsb_mat m : {aproj_H, aproj_HT, aproj_inner_solve})
        if (m) NSB_REQUIRE(m->rows == n_proj, "nsb_simple_create: the condensation operators of aproj must have %lld rows", (long long)n_proj);
    if (pre_mstar) NSB_REQUIRE(pre_mstar->n == n_x, "nsb_simple_create: mstar preconditioner size mismatch");
    if (pre_aproj) NSB_REQUIRE(pre_aproj->n == n_proj, "nsb_simple_create: aproj preconditioner size mismatch");
    NSB_REQUIRE(mstar_maxsteps >= 0 && proj_maxsteps >= 0, "nsb_simple_create: negative step cap");
    Context& c = ctx->c;
    NSB_CUDA(cudaSetDevice(c.device));
    std::vector<int32_t> hmap((size_t)n_vel);
    NSB_CUDA(cudaMemcpy(hmap.data(), map_ind, sizeof(int32_t) * (size_t)n_vel, cudaMemcpyDefault));
    int32_t mx = -1, mn = 0;
    for (int32_t v : hmap) {
        mx = v > mx ? v : mx;
        mn = v < mn ? v : mn;
    }
    NSB_REQUIRE(mn >= 0 && mx < aproj->cols, "nsb_simple_create: map_ind must index the projection vector [owned | halo of aproj]");
    nsb_simple_s* s = new nsb_simple_s();
    s->ctx = &c;
    s->a = a; s->mstar = mstar; s->mH = mstar_H; s->mHT = mstar_HT; s->m_inner = mstar_inner_solve;
    s->bproj = bproj; s->aproj = aproj; s->pH = aproj_H; s->pHT = aproj_HT; s->p_inner = aproj_inner_solve;
    s->pre_m = pre_mstar; s->pre_p = pre_aproj;
    s->n_x = n_x; s->n_vel = n_vel; s->n_proj = n_proj; s->p_begin = p_begin; s->p_end = p_end;
    s->dt = timestep; s->prec_m = mstar_precision; s->prec_p = proj_precision; s->max_m = mstar_maxsteps; s->max_p = proj_maxsteps;
    s->map_needs_halo = mx >= n_proj;
    int r = s->mapV.alloc(n_vel > 0 ? n_vel : 1);
    if (r == NSB_OK && n_vel > 0 && cudaMemcpy(s->mapV.p, hmap.data(), sizeof(int32_t) * (size_t)n_vel, cudaMemcpyHostToDevice) != cudaSuccess) {
        set_error("nsb_simple_create: map upload failed");
        r = NSB_ERR_CUDA;
    }
    if (r == NSB_OK) r = s->xu.alloc(a->cols);
    if (r == NSB_OK) r = s->temp.alloc(max3(n_x, cols_or(mstar_HT, 0), cols_or(mstar_inner_solve, 0)));
    if (r == NSB_OK) r = s->temp2.alloc(n_x);
    if (r == NSB_OK) r = s->velx.alloc(bproj->cols);
    if (r == NSB_OK) r = s->rp.alloc(max3(n_proj, cols_or(aproj_HT, 0), cols_or(aproj_inner_solve, 0)));
    if (r == NSB_OK) r = s->tmp.alloc(aproj->cols > n_proj ? aproj->cols : n_proj);
    if (r == NSB_OK) r = s->conv_d.alloc(n_x);
    if (r == NSB_OK) r = s->f_d.alloc(n_x);
    // CG work vectors are long-lived here (plain allocations, not the stream-ordered pool); u must also hold the halo tail of H
    auto cg_make = [&](CgBuffers& b, nsb_mat_s* M, nsb_mat_s* H, int maxsteps) -> int {
        const int64_t n = M->rows, nx = M->cols > cols_or(H, 0) ? M->cols : cols_or(H, 0);
        NSB_CHECK(b.u.alloc(nx));
        NSB_CHECK(b.s.alloc(M->cols));
        NSB_CHECK(b.d.alloc(n));
        NSB_CHECK(b.w.alloc(n));
        NSB_CHECK(b.q.alloc(n));
        NSB_CHECK(b.hist.alloc(maxsteps + 2));
        return b.st.alloc(1);
    };
    if (r == NSB_OK) r = cg_make(s->cg_m, mstar, mstar_H, mstar_maxsteps);
    if (r == NSB_OK) r = cg_make(s->cg_p, aproj, aproj_H, proj_maxsteps);
    if (r != NSB_OK) {
        delete s;
        return r;
    }
    *out = s;
    return NSB_OK;
}

extern "C" int nsb_simple_destroy(nsb_simple s) {
    delete s;
    return NSB_OK;
}

extern "C" int nsb_simple_step(nsb_ctx ctx, nsb_simple s, double* gfu, const double* f, const double* conv, double* gfup,
                               int* cg_iterations, double* seconds) {
    NSB_REQUIRE(ctx && s && gfu, "nsb_simple_step: NULL argument");
    Context& c = ctx->c;
    NSB_CUDA(cudaSetDevice(c.device));
    NvtxRange nvtx_call("nsb_simple_step");
    EventPair ev;
    NSB_CHECK(ev.create());
    NSB_CUDA(cudaEventRecord(ev.e0, c.stream));
    // gfu: used in place when it is a device vector and `a` needs no halo tail behind it
    const bool in_place = is_device_ptr(gfu) && s->a->cols == s->n_x;
    double* u = in_place ? gfu : s->xu.p;
    if (!in_place) NSB_CHECK(copy_in(c, u, gfu, s->n_x));
    const double* fd = f;
    if (f && !is_device_ptr(f)) {
        NSB_CHECK(copy_in(c, s->f_d.p, f, s->n_x));
        fd = s->f_d.p;
    }
    const double* cd = conv;
    if (conv && !is_device_ptr(conv)) {
        NSB_CHECK(copy_in(c, s->conv_d.p, conv, s->n_x));
        cd = s->conv_d.p;
    }
    int its[2] = {0, 0};
    double* t2 = nullptr;
    if (!s->mH) {   // the right-hand side of the mstar solve straight into the solver's buffer
        NSB_CHECK(apply(c, s->a, u, SimpleRhsOp{cd, fd, s->cg_m.q.p}));
        NSB_CHECK(extended_inverse(c, s->mstar, s->pre_m, nullptr, nullptr, nullptr, s->cg_m, s->g_m, s->cg_m.q.p, true, nullptr, s->prec_m,
                                   s->max_m, &t2, &its[0]));
    } else {
        NSB_CHECK(apply(c, s->a, u, SimpleRhsOp{cd, fd, s->temp.p}));
        NSB_CHECK(extended_inverse(c, s->mstar, s->pre_m, s->mH, s->mHT, s->m_inner, s->cg_m, s->g_m, s->temp.p, false, s->temp2.p,
                                   s->prec_m, s->max_m, &t2, &its[0]));
    }
    // Project(temp2[0:V.ndof]): the velocity part needs room for bproj's halo tail behind it
    double* vel = t2;
    const int64_t t2_cap = (t2 == s->cg_m.u.p) ? s->cg_m.u.n : s->temp2.n;
    if (s->n_vel != s->n_x || t2_cap < s->bproj->cols) {
        NSB_CHECK(vec_lincomb(c, s->velx.p, 1.0, t2, 0, nullptr, 0, nullptr, s->n_vel));
        vel = s->velx.p;
    }
    double* tmp = nullptr;
    NSB_CHECK(project_core(c, s, vel, &its[1], &tmp));
    if (gfup) NSB_CHECK(copy_out(c, gfup, tmp + s->p_begin, s->p_end - s->p_begin));   // :442
    k_simple_update<<<vec_grid(s->n_x, c.num_sms), kBlock, 0, c.stream>>>(u, t2, tmp, s->mapV.p, s->n_vel, s->n_x, s->dt);
    c.launches++;
    NSB_CUDA(cudaGetLastError());
    if (!in_place) NSB_CHECK(copy_out(c, gfu, u, s->n_x));
    NSB_CUDA(cudaEventRecord(ev.e1, c.stream));
    NSB_CUDA(cudaStreamSynchronize(c.stream));
    float ms = 0.f;
    cudaEventElapsedTime(&ms, ev.e0, ev.e1);
    if (cg_iterations) {
        cg_iterations[0] = its[0];
        cg_iterations[1] = its[1];
    }
    if (seconds) *seconds = (double)ms * 1e-3;
    return NSB_OK;
}

extern "C" int nsb_simple_project(nsb_ctx ctx, nsb_simple s, double* vel, double* gfup, int* cg_iterations) {
    NSB_REQUIRE(ctx && s && vel, "nsb_simple_project: NULL argument");
    Context& c = ctx->c;
    NSB_CUDA(cudaSetDevice(c.device));
    NSB_CHECK(copy_in(c, s->velx.p, vel, s->n_vel));
    int its = 0;
    double* tmp = nullptr;
    NSB_CHECK(project_core(c, s, s->velx.p, &its, &tmp));
    if (gfup) NSB_CHECK(copy_out(c, gfup, tmp + s->p_begin, s->p_end - s->p_begin));
    k_simple_correct<<<vec_grid(s->n_vel > 0 ? s->n_vel : 1, c.num_sms), kBlock, 0, c.stream>>>(s->velx.p, tmp, s->mapV.p, s->n_vel);
    c.launches++;
    NSB_CUDA(cudaGetLastError());
    NSB_CHECK(copy_out(c, vel, s->velx.p, s->n_vel));
    NSB_CUDA(cudaStreamSynchronize(c.stream));
    if (cg_iterations) *cg_iterations = its;
    return NSB_OK;
}
