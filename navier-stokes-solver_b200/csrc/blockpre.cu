// blockpre.cu -- block smoother: a.mat.CreateBlockSmoother(blocks) with its three uses in MypreA
// (templates/NavierStokesSIMPLE_iterative.py:360-383):
//   jacobi * x            additive block Jacobi            y   = sum_b P_b^T A_bb^-1 P_b x            (:383)
//   jacobi.Smooth(y, x)   forward block Gauss-Seidel       y_b += A_bb^-1 (x - A y)_b, b ascending    (:377)
//   jacobi.SmoothBack     the same sweep, b descending                                                  (:381)
// Blocks are disjoint dof lists (the facet blocks of :360-362), handed over colour by colour: blocks of one
// colour are not coupled through A, so a colour is one parallel launch and the sweep equals the sequential
// sweep in that block order.  Dense inverses are computed once on the device (Gauss-Jordan in shared memory).
// Row-partitioned: the smoother belongs to this rank's rows of A (columns [owned | halo]); blocks list OWNED dofs
// only (a block is never cut by the partition) under a colouring of the GLOBAL block graph, every rank passing all colours (some may
// be empty on a rank).  Block Jacobi is then purely local.  A Gauss-Seidel sweep reads y at halo columns in its residual rows, so it
// runs on a copy of y with a halo tail that is exchanged before every colour -- the values the neighbours wrote in the colours before
// -- which is the sequential sweep in the global colour-major block order.  The auxiliary-space composite exchanges the input of each
// of its SpMVs (smoother matrix, transform^T over the velocity space, transform over the auxiliary space) the same way.
#include "cond.cuh"

namespace nsb {

constexpr int kMaxBlockDofs = 64;

// one CUDA block per smoother block: gather A_bb from CSR, invert
__global__ void __launch_bounds__(64) k_block_invert(CsrView A, const int64_t* __restrict__ bptr, const int* __restrict__ bdofs,
                                                     const int64_t* __restrict__ iptr, double* __restrict__ inv, int64_t nblocks,
                                                     int* __restrict__ bad) {
    __shared__ double M[kMaxBlockDofs][kMaxBlockDofs + 1];
    __shared__ int dofs[kMaxBlockDofs];
    for (int64_t b = blockIdx.x; b < nblocks; b += gridDim.x) {
        const int64_t d0 = bptr[b];
        const int s = (int)(bptr[b + 1] - d0);
        const int t = threadIdx.x;
        __syncthreads();
        if (t < s) dofs[t] = bdofs[d0 + t];
        __syncthreads();
        if (t < s) {
            for (int j = 0; j < s; ++j) M[t][j] = 0.0;
            const int r = dofs[t];
            for (int p = A.rowptr[r]; p < A.rowptr[r + 1]; ++p) {
                const int col = A.colidx[p];
                for (int j = 0; j < s; ++j)
                    if (dofs[j] == col) M[t][j] += A.values[p];
            }
        }
        __syncthreads();
        for (int k = 0; k < s; ++k) {           // in-place Gauss-Jordan without pivoting (A_bb is SPD on free dofs)
            const double piv = M[k][k];
            if (t == 0 && !(fabs(piv) > 0.0)) atomicExch(bad, 1);
            __syncthreads();
            if (t < s) M[k][t] = (t == k) ? 1.0 / piv : M[k][t] / piv;    // thread t owns column t of the pivot row
            __syncthreads();
            if (t < s && t != k) {                                          // thread t owns row t
                const double f = M[t][k];
                for (int j = 0; j < s; ++j) M[t][j] = (j == k) ? -f * M[k][k] : M[t][j] - f * M[k][j];
            }
            __syncthreads();
        }
        if (t < s)
            for (int j = 0; j < s; ++j) inv[iptr[b] + (int64_t)t * s + j] = M[t][j];
    }
}

// additive apply / one colour of a Gauss-Seidel sweep: one warp per block
template <bool GS>
__global__ void __launch_bounds__(kBlock) k_block_apply(CsrView A, const int64_t* __restrict__ bptr, const int* __restrict__ bdofs,
                                                        const int64_t* __restrict__ iptr, const double* __restrict__ inv,
                                                        int64_t b0, int64_t b1, const double* __restrict__ x, double* y) {
    __shared__ double sr[kBlock / 32][kMaxBlockDofs];
    const int lane = threadIdx.x & 31, w = threadIdx.x >> 5;
    const int64_t warp = ((int64_t)blockIdx.x * blockDim.x + threadIdx.x) >> 5;
    const int64_t nwarps = ((int64_t)gridDim.x * blockDim.x) >> 5;
    for (int64_t b = b0 + warp; b < b1; b += nwarps) {
        const int64_t d0 = bptr[b];
        const int s = (int)(bptr[b + 1] - d0);
        __syncwarp();
        for (int i = lane; i < s; i += 32) {
            const int r = bdofs[d0 + i];
            double v = x[r];
            if (GS) {                                         // residual row: (x - A y)_r
                for (int p = A.rowptr[r]; p < A.rowptr[r + 1]; ++p) v -= A.values[p] * y[A.colidx[p]];
            }
            sr[w][i] = v;
        }
        __syncwarp();
        const double* Ib = inv + iptr[b];
        for (int i = lane; i < s; i += 32) {
            double acc = 0.0;
            for (int j = 0; j < s; ++j) acc = fma(Ib[(int64_t)i * s + j], sr[w][j], acc);
            const int r = bdofs[d0 + i];
            y[r] = GS ? y[r] + acc : acc;
        }
    }
}

static int block_grid(Context& c, int64_t nb) {
    int64_t want = (nb + (kBlock / 32) - 1) / (kBlock / 32);
    if (want < 1) want = 1;
    const int64_t cap = (int64_t)c.num_sms * 8;
    return (int)(want < cap ? want : cap);
}

// y = jacobi * x
static int block_jacobi(Context& c, nsb_pre_s* p, const double* x, double* y) {
    NSB_CHECK(vec_set(c, y, 0.0, p->n));
    if (p->nblocks == 0) return NSB_OK;
    k_block_apply<false><<<block_grid(c, p->nblocks), kBlock, 0, c.stream>>>(view_of(p->mat), p->block_ptr.p, p->block_dofs.p,
                                                                            p->inv_ptr.p, p->inv.p, 0, p->nblocks, x, y);
    c.launches++;
    NSB_CUDA(cudaGetLastError());
    return NSB_OK;
}

// jacobi.Smooth(y, x) / jacobi.SmoothBack(y, x): colours ascending / descending
int block_gs_sweep(Context& c, nsb_pre_s* p, const double* x, double* y, bool backward) {
    nsb_halo_s* plan = p->mat->halo;
    double* yy = y;
    if (plan) {   // row-partitioned: sweep on a copy of y that has a halo tail (y itself holds the owned entries only)
        NSB_CHECK(vec_lincomb(c, p->ybuf.p, 1.0, y, 0, nullptr, 0, nullptr, p->n));
        yy = p->ybuf.p;
    }
    for (int k = 0; k < p->ncolors; ++k) {
        const int col = backward ? p->ncolors - 1 - k : k;
        const int64_t b0 = p->color_ptr[col], b1 = p->color_ptr[col + 1];
        // what the neighbours wrote in the colours before; every rank takes part, also one without blocks of this colour
        if (plan) NSB_CHECK(halo_exchange(c, plan, yy));
        if (b1 <= b0) continue;
        k_block_apply<true><<<block_grid(c, b1 - b0), kBlock, 0, c.stream>>>(view_of(p->mat), p->block_ptr.p, p->block_dofs.p,
                                                                            p->inv_ptr.p, p->inv.p, b0, b1, x, yy);
        c.launches++;
    }
    NSB_CUDA(cudaGetLastError());
    if (plan) NSB_CHECK(vec_lincomb(c, y, 1.0, yy, 0, nullptr, 0, nullptr, p->n));
    return NSB_OK;
}

// MypreA.Mult (:374-383).  x, y: length n, not aliased.
static int auxspace_apply(Context& c, nsb_pre_s* p, const double* x, double* y) {
    nsb_pre_s* sm = p->smoother;
    const double* coarse_in = x;
    if (p->gs) {
        NSB_CHECK(vec_set(c, y, 0.0, p->n));                                                  // y[:] = 0
        NSB_CHECK(block_gs_sweep(c, sm, x, y, false));                                        // jacobi.Smooth(y, x)
        const double* yin = y;
        if (sm->mat->halo) {   // row-partitioned: mat * y reads y's halo tail
            NSB_CHECK(vec_lincomb(c, sm->ybuf.p, 1.0, y, 0, nullptr, 0, nullptr, p->n));
            NSB_CHECK(halo_exchange(c, sm->mat->halo, sm->ybuf.p));
            yin = sm->ybuf.p;
        }
        NSB_CHECK(spmv_launch<0>(c, sm->mat, yin, SubInOp{x, p->aux_t.p}));                   // temp = x - mat * y
        coarse_in = p->aux_t.p;
    } else {
        NSB_CHECK(block_jacobi(c, sm, x, y));                                                 // jacobi * x
    }
    if (p->transform_t->halo) {   // row-partitioned: transform.T reads the velocity halo of its input (aux_t has the tail)
        if (coarse_in != p->aux_t.p) NSB_CHECK(vec_lincomb(c, p->aux_t.p, 1.0, x, 0, nullptr, 0, nullptr, p->n));
        coarse_in = p->aux_t.p;
        NSB_CHECK(halo_exchange(c, p->transform_t->halo, p->aux_t.p));
    }
    NSB_CHECK(spmv_launch<0>(c, p->transform_t, coarse_in, PlainRowOp{p->aux_c.p}));          // transform.T * temp
    NSB_CHECK(pre_apply_device(c, p->coarse, p->aux_c.p, p->aux_z.p));                        // preAh1 * ..
    NSB_CHECK(halo_exchange(c, p->transform->halo, p->aux_z.p));                              // (no-op on one GPU)
    NSB_CHECK(spmv_launch<0>(c, p->transform, p->aux_z.p, AddInOp{y, y}));                    // y += transform * ..
    if (p->gs) NSB_CHECK(block_gs_sweep(c, sm, x, y, true));                                  // jacobi.SmoothBack(y, x)
    return NSB_OK;
}

int pre_apply_device(Context& c, nsb_pre_s* p, const double* x, double* y) {
    if (p->kind == 0) return vec_diagmul(c, y, 1.0, p->d.p, x, p->n);
    if (p->kind == 2) return auxspace_apply(c, p, x, y);
    if (p->kind == 3) {
        NSB_CHECK(pre_apply_device(c, p->blk0, x, y));
        return pre_apply_device(c, p->blk1, x + p->blk0->n, y + p->blk0->n);
    }
    if (!p->gs) return block_jacobi(c, p, x, y);
    // MypreA.Mult with GS = True and no auxiliary-space term (:375-381): y = 0; Smooth(y, x); SmoothBack(y, x)
    NSB_CHECK(vec_set(c, y, 0.0, p->n));
    NSB_CHECK(block_gs_sweep(c, p, x, y, false));
    return block_gs_sweep(c, p, x, y, true);
}

}  // namespace nsb

using namespace nsb;

extern "C" int nsb_pre_create_blocks(nsb_ctx ctx, nsb_mat A, int64_t nblocks, const int64_t* block_ptr, const int32_t* block_dofs,
                                     int ncolors, const int64_t* color_ptr, int gauss_seidel, nsb_pre* out) {
    NSB_REQUIRE(ctx && A && out && block_ptr && nblocks >= 0, "nsb_pre_create_blocks: bad argument");
    NSB_REQUIRE(A->halo ? (A->rows == A->halo->n_owned && A->cols == A->halo->n_owned + A->halo->n_halo) : A->rows == A->cols,
                "nsb_pre_create_blocks: square operator needed (row-partitioned: owned rows, [owned | halo] columns)");
    NSB_REQUIRE(ncolors >= 0 && (ncolors == 0 || color_ptr), "nsb_pre_create_blocks: colour table missing");
    NSB_REQUIRE(!gauss_seidel || ncolors > 0, "nsb_pre_create_blocks: Gauss-Seidel sweeps need the colouring");
    Context& c = ctx->c;
    NSB_CUDA(cudaSetDevice(c.device));
    const int64_t ndofs = block_ptr[nblocks];
    NSB_REQUIRE(block_ptr[0] == 0 && ndofs >= 0 && (ndofs == 0 || block_dofs), "nsb_pre_create_blocks: bad block table");
    std::vector<int64_t> iptr(nblocks + 1, 0);
    int maxb = 0;
    for (int64_t b = 0; b < nblocks; ++b) {
        const int64_t s = block_ptr[b + 1] - block_ptr[b];
        NSB_REQUIRE(s >= 0 && s <= kMaxBlockDofs, "nsb_pre_create_blocks: block %lld has %lld dofs (limit %d)", (long long)b,
                    (long long)s, kMaxBlockDofs);
        iptr[b + 1] = iptr[b] + s * s;
        if ((int)s > maxb) maxb = (int)s;
    }
    if (ncolors > 0) NSB_REQUIRE(color_ptr[0] == 0 && color_ptr[ncolors] == nblocks, "nsb_pre_create_blocks: colours must cover all blocks");
    nsb_pre_s* p = new nsb_pre_s();
    p->ctx = &c;
    p->n = A->rows;
    p->kind = 1;
    p->mat = A;
    p->nblocks = nblocks;
    p->ncolors = ncolors;
    p->max_block = maxb;
    p->gs = gauss_seidel ? 1 : 0;
    if (ncolors > 0) p->color_ptr.assign(color_ptr, color_ptr + ncolors + 1);
    DevBuf<int> bad;
    int r = p->block_ptr.alloc(nblocks + 1);
    if (r == NSB_OK) r = p->block_dofs.alloc(ndofs > 0 ? ndofs : 1);
    if (r == NSB_OK) r = p->inv_ptr.alloc(nblocks + 1);
    if (r == NSB_OK) r = p->inv.alloc(iptr[nblocks] > 0 ? iptr[nblocks] : 1);
    if (r == NSB_OK) r = bad.alloc(1);
    if (r == NSB_OK && A->halo) r = p->ybuf.alloc(A->cols);   // Gauss-Seidel sweeps of a row-partitioned smoother (block_gs_sweep)
    cudaError_t e = cudaSuccess;
    if (r == NSB_OK) {
        e = cudaMemcpy(p->block_ptr.p, block_ptr, sizeof(int64_t) * (nblocks + 1), cudaMemcpyDefault);
        if (e == cudaSuccess && ndofs > 0) e = cudaMemcpy(p->block_dofs.p, block_dofs, sizeof(int32_t) * ndofs, cudaMemcpyDefault);
        if (e == cudaSuccess) e = cudaMemcpy(p->inv_ptr.p, iptr.data(), sizeof(int64_t) * (nblocks + 1), cudaMemcpyHostToDevice);
        if (e == cudaSuccess) e = cudaMemset(bad.p, 0, sizeof(int));
    }
    int hbad = 0;
    if (r == NSB_OK && e == cudaSuccess && nblocks > 0) {
        const int64_t g = nblocks < (int64_t)c.num_sms * 16 ? nblocks : (int64_t)c.num_sms * 16;
        k_block_invert<<<(int)g, 64, 0, c.stream>>>(view_of(A), p->block_ptr.p, p->block_dofs.p, p->inv_ptr.p, p->inv.p, nblocks, bad.p);
        c.launches++;
        e = cudaStreamSynchronize(c.stream);
        if (e == cudaSuccess) e = cudaMemcpy(&hbad, bad.p, sizeof(int), cudaMemcpyDeviceToHost);
    }
    if (r == NSB_OK && e != cudaSuccess) {
        set_error("nsb_pre_create_blocks: %s", cudaGetErrorString(e));
        r = NSB_ERR_CUDA;
    }
    if (r == NSB_OK && hbad) {
        set_error("nsb_pre_create_blocks: a diagonal block is singular (zero pivot)");
        r = NSB_ERR_BREAKDOWN;
    }
    if (r != NSB_OK) {
        delete p;
        return r;
    }
    *out = p;
    return NSB_OK;
}

extern "C" int nsb_pre_smooth(nsb_ctx ctx, nsb_pre p, const double* x, double* y, int backward) {
    NSB_REQUIRE(ctx && p && x && y, "nsb_pre_smooth: NULL argument");
    NSB_REQUIRE(p->kind == 1 && p->ncolors > 0, "nsb_pre_smooth: needs a coloured block smoother");
    Context& c = ctx->c;
    NSB_CUDA(cudaSetDevice(c.device));
    DevBuf<double> dx, dy;
    NSB_CHECK(dx.allocs(c, p->n));
    NSB_CHECK(dy.allocs(c, p->n));
    NSB_CHECK(copy_in(c, dx.p, x, p->n));
    NSB_CHECK(copy_in(c, dy.p, y, p->n));
    NSB_CHECK(block_gs_sweep(c, p, dx.p, dy.p, backward != 0));
    NSB_CHECK(copy_out(c, y, dy.p, p->n));
    NSB_CUDA(cudaStreamSynchronize(c.stream));
    return NSB_OK;
}

extern "C" int nsb_pre_create_auxspace(nsb_ctx ctx, nsb_pre smoother, nsb_mat transform, nsb_mat transform_t, nsb_pre coarse,
                                       int gauss_seidel, nsb_pre* out) {
    NSB_REQUIRE(ctx && smoother && transform && transform_t && coarse && out, "nsb_pre_create_auxspace: bad argument");
    NSB_REQUIRE(smoother->kind == 1, "nsb_pre_create_auxspace: the smoother must come from nsb_pre_create_blocks");
    NSB_REQUIRE(!gauss_seidel || smoother->ncolors > 0, "nsb_pre_create_auxspace: the multiplicative variant needs a coloured smoother");
    const int64_t n = smoother->n, nc = coarse->n;
    // one GPU: transform is n x nc, transform_t nc x n.  Row-partitioned: the owned rows of each, transform's columns numbered over a
    // halo plan of the auxiliary space (n_owned == nc), transform_t's over a halo plan of the velocity space (n_owned == n)
    NSB_REQUIRE(transform->rows == n && transform_t->rows == nc &&
                    (transform->halo ? transform->halo->n_owned == nc : transform->cols == nc) &&
                    (transform_t->halo ? transform_t->halo->n_owned == n : transform_t->cols == n),
                "nsb_pre_create_auxspace: transform must be %lld x %lld and transform_t its transpose", (long long)n, (long long)nc);
    Context& c = ctx->c;
    NSB_CUDA(cudaSetDevice(c.device));
    nsb_pre_s* p = new nsb_pre_s();
    p->ctx = &c;
    p->n = n;
    p->kind = 2;
    p->gs = gauss_seidel ? 1 : 0;
    p->smoother = smoother;
    p->coarse = coarse;
    p->transform = transform;
    p->transform_t = transform_t;
    int rc = p->aux_t.alloc(transform_t->cols > n ? transform_t->cols : n);   // SpMV input of transform_t: owned + halo entries
    if (rc == NSB_OK) rc = p->aux_c.alloc(nc);
    if (rc == NSB_OK) rc = p->aux_z.alloc(transform->cols > nc ? transform->cols : nc);   // SpMV input of transform
    if (rc != NSB_OK) {
        delete p;
        return rc;
    }
    *out = p;
    return NSB_OK;
}
