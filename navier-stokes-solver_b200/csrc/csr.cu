// csr.cu -- operator handles: CSR upload, device transpose, saddle-point composition, diagonal,
// plain SpMV entry points and the diagonal preconditioner handles.
#include <cub/cub.cuh>
#include <stdlib.h>

#include "spmv.cuh"

namespace nsb {

static int pick_tpr(int64_t rows, int64_t nnz) {
    const char* env = getenv("NSB_TPR");
    if (env) {
        int v = atoi(env);
        if (v == 2 || v == 4 || v == 8 || v == 16 || v == 32) return v;
    }
    double mean = rows > 0 ? (double)nnz / (double)rows : 0.0;
    if (mean <= 6.0) return 2;
    if (mean <= 12.0) return 4;
    if (mean <= 40.0) return 8;
    if (mean <= 96.0) return 16;
    return 32;
}

// allocate the three arrays; the two streams are padded to an even length + 2 (spmv.cuh)
static int mat_alloc(Context& c, int64_t rows, int64_t cols, int64_t nnz, nsb_mat_s** out) {
    NSB_REQUIRE(rows >= 0 && cols >= 0 && nnz >= 0, "matrix with negative size");
    NSB_REQUIRE(nnz < 2147483000LL, "nnz = %lld exceeds the int32 row-pointer range of this build", (long long)nnz);
    NSB_REQUIRE(rows < 2147483000LL && cols < 2147483000LL, "matrix dimension exceeds int32 index range");
    nsb_mat_s* m = new nsb_mat_s();
    m->ctx = &c;
    m->rows = rows;
    m->cols = cols;
    m->nnz = nnz;
    int64_t padded = ((nnz + 1) & ~1LL) + 2;
    int r = m->rowptr.alloc(rows + 1);
    if (r == NSB_OK) r = m->colidx.alloc(padded);
    if (r == NSB_OK) r = m->values.alloc(padded);
    if (r == NSB_OK) {
        cudaError_t e = cudaMemsetAsync(m->colidx.p + (padded - 4 > 0 ? padded - 4 : 0), 0, sizeof(int32_t) * (padded < 4 ? padded : 4), c.stream);
        if (e == cudaSuccess) e = cudaMemsetAsync(m->values.p + (padded - 4 > 0 ? padded - 4 : 0), 0, sizeof(double) * (padded < 4 ? padded : 4), c.stream);
        if (e != cudaSuccess) {
            set_error("memset failed: %s", cudaGetErrorString(e));
            r = NSB_ERR_CUDA;
        }
    }
    if (r != NSB_OK) {
        delete m;
        return r;
    }
    m->threads_per_row = pick_tpr(rows, nnz);
    *out = m;
    return NSB_OK;
}

__global__ void k_rowptr64_to_32(const int64_t* __restrict__ in, int* __restrict__ out, int64_t n) {
    for (int64_t i = (int64_t)blockIdx.x * blockDim.x + threadIdx.x; i < n; i += (int64_t)gridDim.x * blockDim.x) out[i] = (int)in[i];
}

// ---- transpose -------------------------------------------------------------------------------
__global__ void k_transpose_keys(CsrView A, int64_t cols_shift_unused, long long* __restrict__ keys, int* __restrict__ perm) {
    // one warp per row: key = (col << 32) | row   -> sorting yields the CSR of A^T with sorted columns
    const int lane = threadIdx.x & 31;
    const int64_t warp = ((int64_t)blockIdx.x * blockDim.x + threadIdx.x) >> 5;
    const int64_t nwarps = ((int64_t)gridDim.x * blockDim.x) >> 5;
    for (int64_t r = warp; r < A.rows; r += nwarps) {
        const int p0 = A.rowptr[r], p1 = A.rowptr[r + 1];
        for (int p = p0 + lane; p < p1; p += 32) {
            keys[p] = ((long long)A.colidx[p] << 32) | (long long)r;
            perm[p] = p;
        }
    }
}

__global__ void k_transpose_fill(const long long* __restrict__ keys, const int* __restrict__ perm,
                                 const double* __restrict__ vin, int* __restrict__ colidx, double* __restrict__ vout,
                                 int* __restrict__ counts, int64_t nnz) {
    for (int64_t i = (int64_t)blockIdx.x * blockDim.x + threadIdx.x; i < nnz; i += (int64_t)gridDim.x * blockDim.x) {
        const long long k = keys[i];
        colidx[i] = (int)(k & 0xffffffffLL);
        vout[i] = vin[perm[i]];
        atomicAdd(&counts[(int)(k >> 32) + 1], 1);  // integer histogram: order-independent
    }
}

// ---- saddle composition ------------------------------------------------------------------------
__global__ void k_saddle_rowptr(CsrView A, CsrView BT, CsrView B, CsrView C, bool hasC, int* __restrict__ rowptr) {
    const int64_t n_u = A.rows, n_p = B.rows;
    const int top = A.rowptr[n_u] + BT.rowptr[n_u];
    for (int64_t i = (int64_t)blockIdx.x * blockDim.x + threadIdx.x; i <= n_u + n_p; i += (int64_t)gridDim.x * blockDim.x) {
        if (i <= n_u) rowptr[i] = A.rowptr[i] + BT.rowptr[i];
        else {
            const int64_t j = i - n_u;
            rowptr[i] = top + B.rowptr[j] + (hasC ? C.rowptr[j] : 0);
        }
    }
}

__global__ void k_saddle_fill(CsrView A, CsrView BT, CsrView B, CsrView C, bool hasC, const int* __restrict__ rowptr,
                              int* __restrict__ colidx, double* __restrict__ values) {
    const int lane = threadIdx.x & 31;
    const int64_t warp = ((int64_t)blockIdx.x * blockDim.x + threadIdx.x) >> 5;
    const int64_t nwarps = ((int64_t)gridDim.x * blockDim.x) >> 5;
    const int64_t n_u = A.rows, n_p = B.rows;
    for (int64_t i = warp; i < n_u + n_p; i += nwarps) {
        const bool topRow = i < n_u;
        const CsrView& L = topRow ? A : B;
        const CsrView& R = topRow ? BT : C;
        const int64_t j = topRow ? i : i - n_u;
        int dst = rowptr[i];
        const int l0 = L.rowptr[j], l1 = L.rowptr[j + 1];
        for (int p = l0 + lane; p < l1; p += 32) {
            colidx[dst + (p - l0)] = L.colidx[p];
            values[dst + (p - l0)] = L.values[p];
        }
        dst += l1 - l0;
        if (topRow || hasC) {
            const int r0 = R.rowptr[j], r1 = R.rowptr[j + 1];
            for (int p = r0 + lane; p < r1; p += 32) {
                colidx[dst + (p - r0)] = R.colidx[p] + (int)n_u;
                values[dst + (p - r0)] = R.values[p];
            }
        }
    }
}

__global__ void k_diagonal(CsrView A, double* __restrict__ d) {
    for (int64_t r = (int64_t)blockIdx.x * blockDim.x + threadIdx.x; r < A.rows; r += (int64_t)gridDim.x * blockDim.x) {
        double v = 0.0;
        for (int p = A.rowptr[r]; p < A.rowptr[r + 1]; ++p)
            if (A.colidx[p] == r) v += A.values[p];
        d[r] = v;
    }
}

__global__ void k_jacobi_invert(double* __restrict__ d, const unsigned char* __restrict__ free_mask, int64_t n) {
    for (int64_t i = (int64_t)blockIdx.x * blockDim.x + threadIdx.x; i < n; i += (int64_t)gridDim.x * blockDim.x) {
        const double v = d[i];
        const bool on = (free_mask == nullptr || free_mask[i] != 0) && v != 0.0;
        d[i] = on ? 1.0 / v : 0.0;
    }
}

int spmv_plain(Context& c, const nsb_mat_s* m, const double* x, double* y) {
    PlainRowOp op{y};
    return spmv_launch<0>(c, m, x, op);
}

}  // namespace nsb

using namespace nsb;

extern "C" {

int nsb_csr_create(nsb_ctx ctx, int64_t rows, int64_t cols, int64_t nnz, const void* rowptr, int rowptr_bits,
                   const int32_t* colidx, const double* values, nsb_mat* out) {
    NSB_REQUIRE(ctx && out, "nsb_csr_create: NULL argument");
    NSB_REQUIRE(rowptr_bits == 32 || rowptr_bits == 64, "nsb_csr_create: rowptr_bits must be 32 or 64");
    NSB_REQUIRE(rowptr && (nnz == 0 || (colidx && values)), "nsb_csr_create: NULL array");
    Context& c = ctx->c;
    NSB_CUDA(cudaSetDevice(c.device));
    cudaGetLastError();   // a non-sticky error some earlier caller never collected must not be blamed on the cub calls below
    nsb_mat_s* m = nullptr;
    NSB_CHECK(mat_alloc(c, rows, cols, nnz, &m));
    int r = NSB_OK;
    cudaError_t e = cudaSuccess;
    if (rowptr_bits == 32) {
        e = cudaMemcpyAsync(m->rowptr.p, rowptr, sizeof(int32_t) * (size_t)(rows + 1), cudaMemcpyDefault, c.stream);
    } else {
        DevBuf<int64_t> tmp;
        r = tmp.alloc(rows + 1);
        if (r == NSB_OK) {
            e = cudaMemcpyAsync(tmp.p, rowptr, sizeof(int64_t) * (size_t)(rows + 1), cudaMemcpyDefault, c.stream);
            k_rowptr64_to_32<<<vec_grid(rows + 1, c.num_sms), kBlock, 0, c.stream>>>(tmp.p, m->rowptr.p, rows + 1);
            if (e == cudaSuccess) e = cudaStreamSynchronize(c.stream);
        }
    }
    if (r == NSB_OK && e == cudaSuccess && nnz > 0) {
        e = cudaMemcpyAsync(m->colidx.p, colidx, sizeof(int32_t) * (size_t)nnz, cudaMemcpyDefault, c.stream);
        if (e == cudaSuccess) e = cudaMemcpyAsync(m->values.p, values, sizeof(double) * (size_t)nnz, cudaMemcpyDefault, c.stream);
    }
    if (e == cudaSuccess) e = cudaStreamSynchronize(c.stream);
    if (r == NSB_OK && e != cudaSuccess) {
        set_error("nsb_csr_create: upload failed: %s", cudaGetErrorString(e));
        r = NSB_ERR_CUDA;
    }
    if (r == NSB_OK) {
        // validate the row-pointer ends (cheap; catches a wrong rowptr_bits)
        int ends[2] = {0, 0};
        cudaMemcpy(&ends[0], m->rowptr.p, sizeof(int), cudaMemcpyDeviceToHost);
        cudaMemcpy(&ends[1], m->rowptr.p + rows, sizeof(int), cudaMemcpyDeviceToHost);
        if (ends[0] != 0 || ends[1] != (int)nnz) {
            set_error("nsb_csr_create: rowptr[0]=%d, rowptr[rows]=%d, expected 0 and %lld", ends[0], ends[1], (long long)nnz);
            r = NSB_ERR_INVALID;
        }
    }
    if (r == NSB_OK) r = build_sell(c, m);
    if (r != NSB_OK) {
        delete m;
        return r;
    }
    *out = m;
    return NSB_OK;
}

int nsb_csr_transpose(nsb_ctx ctx, nsb_mat a, nsb_mat* out) {
    NSB_REQUIRE(ctx && a && out, "nsb_csr_transpose: NULL argument");
    NSB_REQUIRE(a->halo == nullptr, "nsb_csr_transpose: not available for row-partitioned operators (upload the local rows of the transpose)");
    Context& c = ctx->c;
    NSB_CUDA(cudaSetDevice(c.device));
    nsb_mat_s* t = nullptr;
    NSB_CHECK(mat_alloc(c, a->cols, a->rows, a->nnz, &t));
    const int64_t nnz = a->nnz;
    int r = NSB_OK;
    DevBuf<long long> k0, k1;
    DevBuf<int> p0, p1;
    DevBuf<char> tmp;
    size_t tmp_bytes = 0;
    cudaError_t e = cudaMemsetAsync(t->rowptr.p, 0, sizeof(int) * (size_t)(t->rows + 1), c.stream);
    if (nnz > 0 && e == cudaSuccess) {
        if ((r = k0.alloc(nnz)) == NSB_OK && (r = k1.alloc(nnz)) == NSB_OK && (r = p0.alloc(nnz)) == NSB_OK &&
            (r = p1.alloc(nnz)) == NSB_OK) {
            CsrView A = view_of(a);
            k_transpose_keys<<<spmv_grid(a->rows, 32, c.num_sms), kBlock, 0, c.stream>>>(A, 0, k0.p, p0.p);
            cub::DeviceRadixSort::SortPairs(nullptr, tmp_bytes, k0.p, k1.p, p0.p, p1.p, (int)nnz, 0, 64, c.stream);
            if ((r = tmp.alloc((int64_t)tmp_bytes + 16)) == NSB_OK) {
                e = cub::DeviceRadixSort::SortPairs(tmp.p, tmp_bytes, k0.p, k1.p, p0.p, p1.p, (int)nnz, 0, 64, c.stream);
                k_transpose_fill<<<vec_grid(nnz, c.num_sms), kBlock, 0, c.stream>>>(k1.p, p1.p, a->values.p, t->colidx.p,
                                                                                  t->values.p, t->rowptr.p, nnz);
            }
        }
    }
    if (r == NSB_OK && e == cudaSuccess) {
        // counts (shifted by one) -> row pointers
        size_t sb = 0;
        cub::DeviceScan::InclusiveSum(nullptr, sb, t->rowptr.p, t->rowptr.p, (int)(t->rows + 1), c.stream);
        DevBuf<char> stmp;
        if ((r = stmp.alloc((int64_t)sb + 16)) == NSB_OK) {
            e = cub::DeviceScan::InclusiveSum(stmp.p, sb, t->rowptr.p, t->rowptr.p, (int)(t->rows + 1), c.stream);
            if (e == cudaSuccess) e = cudaStreamSynchronize(c.stream);
        }
    }
    if (e == cudaSuccess) e = cudaStreamSynchronize(c.stream);
    if (r == NSB_OK && e != cudaSuccess) {
        set_error("nsb_csr_transpose: %s", cudaGetErrorString(e));
        r = NSB_ERR_CUDA;
    }
    if (r == NSB_OK) r = build_sell(c, t);
    if (r != NSB_OK) {
        delete t;
        return r;
    }
    *out = t;
    return NSB_OK;
}

int nsb_saddle_create(nsb_ctx ctx, nsb_mat A, nsb_mat B, nsb_mat BT, nsb_mat C, nsb_mat* out) {
    NSB_REQUIRE(ctx && A && B && out, "nsb_saddle_create: NULL argument");
    NSB_REQUIRE(A->rows == A->cols, "nsb_saddle_create: A must be square (%lld x %lld)", (long long)A->rows, (long long)A->cols);
    NSB_REQUIRE(B->cols == A->rows, "nsb_saddle_create: B has %lld columns, A has %lld rows", (long long)B->cols, (long long)A->rows);
    NSB_REQUIRE(!A->halo && !B->halo, "nsb_saddle_create: row-partitioned blocks are composed on the host side");
    if (C) NSB_REQUIRE(C->rows == B->rows && C->cols == B->rows, "nsb_saddle_create: C must be n_p x n_p");
    Context& c = ctx->c;
    NSB_CUDA(cudaSetDevice(c.device));
    nsb_mat_s* bt_owned = nullptr;
    if (!BT) {
        NSB_CHECK(nsb_csr_transpose(ctx, B, &bt_owned));
        BT = bt_owned;
    }
    int r = NSB_OK;
    nsb_mat_s* K = nullptr;
    if (BT->rows != A->rows || BT->cols != B->rows) {
        set_error("nsb_saddle_create: BT must be n_u x n_p");
        r = NSB_ERR_INVALID;
    }
    const int64_t n_u = A->rows, n_p = B->rows;
    const int64_t nnz = A->nnz + BT->nnz + B->nnz + (C ? C->nnz : 0);
    if (r == NSB_OK) r = mat_alloc(c, n_u + n_p, n_u + n_p, nnz, &K);
    if (r == NSB_OK) {
        CsrView vA = view_of(A), vBT = view_of(BT), vB = view_of(B), vC = C ? view_of(C) : view_of(B);
        k_saddle_rowptr<<<vec_grid(n_u + n_p + 1, c.num_sms), kBlock, 0, c.stream>>>(vA, vBT, vB, vC, C != nullptr, K->rowptr.p);
        k_saddle_fill<<<spmv_grid(n_u + n_p, 32, c.num_sms), kBlock, 0, c.stream>>>(vA, vBT, vB, vC, C != nullptr, K->rowptr.p,
                                                                                   K->colidx.p, K->values.p);
        cudaError_t e = cudaStreamSynchronize(c.stream);
        if (e != cudaSuccess) {
            set_error("nsb_saddle_create: %s", cudaGetErrorString(e));
            r = NSB_ERR_CUDA;
        }
    }
    delete bt_owned;
    // two row panels: velocity rows [A B^T] (~36 nonzeros) and pressure rows [B C] (~190) get their own SELL-LX shapes
    if (r == NSB_OK) K->panel_splits = {0, n_u, n_u + n_p};
    if (r == NSB_OK) r = build_sell(c, K);
    if (r != NSB_OK) {
        delete K;
        return r;
    }
    K->n_u = n_u;
    K->n_p = n_p;
    *out = K;
    return NSB_OK;
}

int nsb_mat_destroy(nsb_mat m) {
    delete m;
    return NSB_OK;
}

int nsb_mat_info(nsb_mat m, int64_t* rows, int64_t* cols, int64_t* nnz) {
    NSB_REQUIRE(m, "nsb_mat_info: NULL matrix");
    if (rows) *rows = m->rows;
    if (cols) *cols = m->cols;
    if (nnz) *nnz = m->nnz;
    return NSB_OK;
}

int nsb_mat_format(nsb_mat m, int* rows_per_chunk, int* chunks, int* max_distinct, int64_t* total_distinct,
                   int* lanes_per_row, int64_t* stored_nnz, int64_t* column_runs) {
    NSB_REQUIRE(m, "nsb_mat_format: NULL matrix");
    // aggregated over the row panels; rows_per_chunk / lanes_per_row describe the first (largest) panel
    const bool sell = !m->panels.empty();
    int chunks_ = 0, maxd = 0;
    int64_t tot = 0, stored = 0, runs = 0;
    for (const auto& P : m->panels) {
        chunks_ += P->sell_nchunks;
        if (P->sell_max_uniq > maxd) maxd = P->sell_max_uniq;
        tot += P->sell_total_uniq;
        stored += P->sell_stored;
        runs += P->sell_nruns;
    }
    if (rows_per_chunk) *rows_per_chunk = sell ? m->panels[0]->sell_R : 0;
    if (chunks) *chunks = chunks_;
    if (max_distinct) *max_distinct = maxd;
    if (total_distinct) *total_distinct = tot;
    if (lanes_per_row) *lanes_per_row = sell ? m->panels[0]->sell_T : m->threads_per_row;
    if (stored_nnz) *stored_nnz = sell ? stored : m->nnz;
    if (column_runs) *column_runs = sell ? runs : 0;
    return NSB_OK;
}

int nsb_mat_set_row_panels(nsb_ctx ctx, nsb_mat m, int nbounds, const int64_t* bounds) {
    NSB_REQUIRE(ctx && m && bounds && nbounds >= 2, "nsb_mat_set_row_panels: bad argument");
    NSB_REQUIRE(bounds[0] == 0 && bounds[nbounds - 1] == m->rows, "nsb_mat_set_row_panels: bounds must run from 0 to rows");
    for (int k = 1; k < nbounds; ++k) NSB_REQUIRE(bounds[k] >= bounds[k - 1], "nsb_mat_set_row_panels: bounds must ascend");
    Context& c = ctx->c;
    NSB_CUDA(cudaSetDevice(c.device));
    NSB_CUDA(cudaStreamSynchronize(c.stream));
    m->panel_splits.assign(bounds, bounds + nbounds);
    return build_sell(c, m);
}

int nsb_mat_diagonal(nsb_ctx ctx, nsb_mat m, double* out) {
    NSB_REQUIRE(ctx && m && out, "nsb_mat_diagonal: NULL argument");
    Context& c = ctx->c;
    NSB_CUDA(cudaSetDevice(c.device));
    DevBuf<double> d;
    NSB_CHECK(d.alloc(m->rows));
    k_diagonal<<<vec_grid(m->rows, c.num_sms), kBlock, 0, c.stream>>>(view_of(m), d.p);
    NSB_CHECK(copy_out(c, out, d.p, m->rows));
    NSB_CUDA(cudaStreamSynchronize(c.stream));
    return NSB_OK;
}

int nsb_spmv(nsb_ctx ctx, nsb_mat m, const double* x, double* y) {
    NSB_REQUIRE(ctx && m && x && y, "nsb_spmv: NULL argument");
    Context& c = ctx->c;
    NSB_CUDA(cudaSetDevice(c.device));
    const int64_t n_in = m->halo ? m->halo->n_owned : m->cols;
    DevBuf<double> dx, dy;
    NSB_CHECK(dx.allocs(c, m->cols));
    NSB_CHECK(dy.allocs(c, m->rows));
    NSB_CHECK(copy_in(c, dx.p, x, n_in));
    NSB_CHECK(halo_exchange(c, m->halo, dx.p));
    NSB_CHECK(spmv_plain(c, m, dx.p, dy.p));
    NSB_CHECK(copy_out(c, y, dy.p, m->rows));
    NSB_CUDA(cudaStreamSynchronize(c.stream));
    return NSB_OK;
}

int nsb_spmv_bench(nsb_ctx ctx, nsb_mat m, int warmup, int reps, double* ms_per_launch) {
    NSB_REQUIRE(ctx && m && ms_per_launch && reps > 0, "nsb_spmv_bench: bad argument");
    Context& c = ctx->c;
    NSB_CUDA(cudaSetDevice(c.device));
    DevBuf<double> dx, dy;
    NSB_CHECK(dx.allocs(c, m->cols));
    NSB_CHECK(dy.allocs(c, m->rows));
    // x ~ deterministic pattern in (-1, 1)
    NSB_CHECK(vec_set(c, dx.p, 0.5, m->cols));
    cudaEvent_t e0, e1;
    NSB_CUDA(cudaEventCreate(&e0));
    NSB_CUDA(cudaEventCreate(&e1));
    // NSB_BENCH_HALOIN=1 (diagnostic): the same SpMV through the kernel variant that takes halo columns from a landing buffer
    // (p2p.cuh: HaloIn) with an empty halo and the identity chunk order -- what that variant costs by itself, on one GPU
    const char* hb = getenv("NSB_BENCH_HALOIN");
    std::vector<std::unique_ptr<DevBuf<int>>> orders;
    const bool halo_variant = hb && atoi(hb) != 0 && !m->panels.empty() && !m->halo;
    if (halo_variant) {
        for (const auto& P : m->panels) {
            std::vector<int> id((size_t)P->sell_nchunks);
            for (int k = 0; k < P->sell_nchunks; ++k) id[(size_t)k] = k;
            orders.emplace_back(new DevBuf<int>());
            NSB_CHECK(orders.back()->alloc(P->sell_nchunks));
            NSB_CUDA(cudaMemcpy(orders.back()->p, id.data(), sizeof(int) * id.size(), cudaMemcpyHostToDevice));
        }
    }
    auto one = [&]() -> int {
        if (!halo_variant) return spmv_plain(c, m, dx.p, dy.p);
        for (size_t k = 0; k < m->panels.size(); ++k) {
            const HaloIn H{nullptr, (const u64*)c.d_mail, dx.p, m->cols, 0, orders[k]->p, m->panels[k]->sell_nchunks, nullptr,
                           SpinGuard{c.fault, 0}};
            NSB_CHECK((spmv_sell_panel<0, PlainRowOp, HaloIn>(c, m->panels[k].get(), dx.p, PlainRowOp{dy.p}, H)));
        }
        return NSB_OK;
    };
    for (int i = 0; i < warmup; ++i) NSB_CHECK(one());
    NSB_CUDA(cudaEventRecord(e0, c.stream));
    for (int i = 0; i < reps; ++i) NSB_CHECK(one());
    NSB_CUDA(cudaEventRecord(e1, c.stream));
    NSB_CUDA(cudaEventSynchronize(e1));
    float ms = 0.f;
    NSB_CUDA(cudaEventElapsedTime(&ms, e0, e1));
    cudaEventDestroy(e0);
    cudaEventDestroy(e1);
    *ms_per_launch = (double)ms / reps;
    return NSB_OK;
}

// ---- preconditioners -------------------------------------------------------------------------
int nsb_pre_create_diag(nsb_ctx ctx, int64_t n, const double* d, nsb_pre* out) {
    NSB_REQUIRE(ctx && d && out && n >= 0, "nsb_pre_create_diag: bad argument");
    Context& c = ctx->c;
    NSB_CUDA(cudaSetDevice(c.device));
    nsb_pre_s* p = new nsb_pre_s();
    p->ctx = &c;
    p->n = n;
    int r = p->d.alloc(n > 0 ? n : 1);
    if (r == NSB_OK) r = copy_in(c, p->d.p, d, n);
    if (r == NSB_OK && cudaStreamSynchronize(c.stream) != cudaSuccess) {
        set_error("nsb_pre_create_diag: copy failed");
        r = NSB_ERR_CUDA;
    }
    if (r != NSB_OK) {
        delete p;
        return r;
    }
    *out = p;
    return NSB_OK;
}

int nsb_pre_create_jacobi(nsb_ctx ctx, nsb_mat m, const uint8_t* free_mask, nsb_pre* out) {
    NSB_REQUIRE(ctx && m && out, "nsb_pre_create_jacobi: NULL argument");
    Context& c = ctx->c;
    NSB_CUDA(cudaSetDevice(c.device));
    nsb_pre_s* p = new nsb_pre_s();
    p->ctx = &c;
    p->n = m->rows;
    DevBuf<unsigned char> mask;
    int r = p->d.alloc(m->rows > 0 ? m->rows : 1);
    if (r == NSB_OK && free_mask) {
        r = mask.alloc(m->rows);
        if (r == NSB_OK && cudaMemcpyAsync(mask.p, free_mask, (size_t)m->rows, cudaMemcpyDefault, c.stream) != cudaSuccess) {
            set_error("nsb_pre_create_jacobi: mask copy failed");
            r = NSB_ERR_CUDA;
        }
    }
    if (r == NSB_OK) {
        k_diagonal<<<vec_grid(m->rows, c.num_sms), kBlock, 0, c.stream>>>(view_of(m), p->d.p);
        k_jacobi_invert<<<vec_grid(m->rows, c.num_sms), kBlock, 0, c.stream>>>(p->d.p, free_mask ? mask.p : nullptr, m->rows);
        if (cudaStreamSynchronize(c.stream) != cudaSuccess) {
            set_error("nsb_pre_create_jacobi: kernel failed: %s", cudaGetErrorString(cudaGetLastError()));
            r = NSB_ERR_CUDA;
        }
    }
    if (r != NSB_OK) {
        delete p;
        return r;
    }
    *out = p;
    return NSB_OK;
}

int nsb_pre_create_blockdiag2(nsb_ctx ctx, nsb_pre a, nsb_pre b, nsb_pre* out) {
    NSB_REQUIRE(ctx && a && b && out, "nsb_pre_create_blockdiag2: NULL argument");
    Context& c = ctx->c;
    NSB_CUDA(cudaSetDevice(c.device));
    nsb_pre_s* p = new nsb_pre_s();
    p->ctx = &c;
    p->n = a->n + b->n;
    if (a->kind != 0 || b->kind != 0) {   // a block smoother / composite in one of the blocks: applied block by block (borrowed handles)
        p->kind = 3;
        p->blk0 = a;
        p->blk1 = b;
        *out = p;
        return NSB_OK;
    }
    int r = p->d.alloc(p->n > 0 ? p->n : 1);
    if (r == NSB_OK) r = copy_in(c, p->d.p, a->d.p, a->n);
    if (r == NSB_OK) r = copy_in(c, p->d.p + a->n, b->d.p, b->n);
    if (r == NSB_OK && cudaStreamSynchronize(c.stream) != cudaSuccess) {
        set_error("nsb_pre_create_blockdiag2: copy failed");
        r = NSB_ERR_CUDA;
    }
    if (r != NSB_OK) {
        delete p;
        return r;
    }
    *out = p;
    return NSB_OK;
}

int nsb_pre_destroy(nsb_pre p) {
    delete p;
    return NSB_OK;
}

int nsb_pre_apply(nsb_ctx ctx, nsb_pre p, const double* x, double* y) {
    NSB_REQUIRE(ctx && p && x && y, "nsb_pre_apply: NULL argument");
    Context& c = ctx->c;
    NSB_CUDA(cudaSetDevice(c.device));
    DevBuf<double> dx, dy;
    NSB_CHECK(dx.allocs(c, p->n));
    NSB_CHECK(dy.allocs(c, p->n));
    NSB_CHECK(copy_in(c, dx.p, x, p->n));
    NSB_CHECK(pre_apply_device(c, p, dx.p, dy.p));
    NSB_CHECK(copy_out(c, y, dy.p, p->n));
    NSB_CUDA(cudaStreamSynchronize(c.stream));
    return NSB_OK;
}

}  // extern "C"
