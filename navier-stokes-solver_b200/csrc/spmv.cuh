// spmv.cuh -- fp64 CSR SpMV kernel template with fused row epilogues (sm_100a).
//
// Replaces NGSolve's SparseMatrix::MultAdd at every `mat * vec` of the reference loops
// (minres.py:97; bramble_pasciak_cg.py:125,127,135; solvers/bramblepasciak_new.py:206,210,213).
//
// Layout in HBM: rowptr int32[rows+1], colidx int32[nnz(+pad)], values fp64[nnz(+pad)]; both streams are
// padded to an even length so that every lane can issue 128-bit (values) / 64-bit (indices) loads
// at even offsets regardless of where a row starts.  Algorithmic traffic: 12 B per nonzero +
// 4 B per row pointer + x once + y once.
//
// Kernel shape ("CSR-vector"): a group of TPR lanes owns a row, lanes stride over the row two
// nonzeros at a time, partial sums are combined with xor-shuffles (fixed tree => bitwise
// reproducible).  Rows are dealt to groups grid-stride from a fixed persistent grid (multiple of the
// SM count), so the whole grid advances through the matrix as one contiguous front of the two
// streams.  The RowOp functor receives (row, (Mx)_row) in one lane and may write any number of
// fused outputs and contribute to NS grid-reduced sums whose totals go to RowOp::finish() in the
// last block (common.cuh: grid_sum).
#pragma once

#include "common.cuh"

namespace nsb {

__device__ __forceinline__ double2 ld_stream_f64x2(const double* p) {
    double2 r;
    asm volatile("ld.global.nc.L1::no_allocate.v2.f64 {%0, %1}, [%2];" : "=d"(r.x), "=d"(r.y) : "l"(p));
    return r;
}
__device__ __forceinline__ int2 ld_stream_s32x2(const int* p) {
    int2 r;
    asm volatile("ld.global.nc.L1::no_allocate.v2.s32 {%0, %1}, [%2];" : "=r"(r.x), "=r"(r.y) : "l"(p));
    return r;
}

struct CsrView {
    int64_t rows;
    const int* __restrict__ rowptr;
    const int* __restrict__ colidx;
    const double* __restrict__ values;
};

inline CsrView view_of(const nsb_mat_s* m) { return CsrView{m->rows, m->rowptr.p, m->colidx.p, m->values.p}; }

template <int TPR, int NS, class RowOp>
__global__ void __launch_bounds__(kBlock) spmv_csrvec_kernel(CsrView A, const double* __restrict__ x, RowOp op,
                                                             double* partials, unsigned int* ticket) {
    if (op.skip()) return;
    constexpr int NSA = NS > 0 ? NS : 1;
    double acc[NSA];
#pragma unroll
    for (int k = 0; k < NSA; ++k) acc[k] = 0.0;

    const int lane = threadIdx.x & (TPR - 1);
    const int sub = (threadIdx.x & 31) / TPR;  // group index inside the warp
    const int64_t wgroup0 = ((int64_t)blockIdx.x * blockDim.x + (threadIdx.x & ~31)) / TPR;
    const int64_t ngroups = ((int64_t)gridDim.x * blockDim.x) / TPR;

    // the row loop is warp-uniform (all groups of a warp iterate together) so that the full-mask
    // shuffles below are always executed by all 32 lanes
    for (int64_t rb = wgroup0; rb < A.rows; rb += ngroups) {
        const int64_t r = rb + sub;
        const bool valid = r < A.rows;
        const int p0 = valid ? __ldg(A.rowptr + r) : 0;
        const int p1 = valid ? __ldg(A.rowptr + r + 1) : 0;
        double s0 = 0.0, s1 = 0.0;
        int p = (p0 & ~1) + 2 * lane;
        // two independent pair-loads in flight per lane
        for (; p + 2 * TPR < p1; p += 4 * TPR) {
            const double2 va = ld_stream_f64x2(A.values + p);
            const int2 ca = ld_stream_s32x2(A.colidx + p);
            const double2 vb = ld_stream_f64x2(A.values + p + 2 * TPR);
            const int2 cb = ld_stream_s32x2(A.colidx + p + 2 * TPR);
            const double xa0 = (p >= p0) ? __ldg(x + ca.x) : 0.0;
            const double xa1 = __ldg(x + ca.y);   // p + 1 < p1 holds here
            const double xb0 = __ldg(x + cb.x);
            const double xb1 = (p + 2 * TPR + 1 < p1) ? __ldg(x + cb.y) : 0.0;
            s0 = fma(va.x, xa0, s0);
            s1 = fma(va.y, xa1, s1);
            s0 = fma(vb.x, xb0, s0);
            s1 = fma(vb.y, xb1, s1);
        }
        if (p < p1) {
            const double2 va = ld_stream_f64x2(A.values + p);
            const int2 ca = ld_stream_s32x2(A.colidx + p);
            const double xa0 = (p >= p0) ? __ldg(x + ca.x) : 0.0;
            const double xa1 = (p + 1 < p1) ? __ldg(x + ca.y) : 0.0;
            s0 = fma(va.x, xa0, s0);
            s1 = fma(va.y, xa1, s1);
        }
        double s = s0 + s1;
#pragma unroll
        for (int o = TPR / 2; o > 0; o >>= 1) s += __shfl_xor_sync(0xffffffffu, s, o);
        if (lane == 0 && valid) op.row(r, s, acc);
    }
    if constexpr (NS > 0) {
        if (grid_sum<NSA>(acc, partials, ticket)) {
            if (threadIdx.x == 0) op.finish(acc);
        }
    }
}

inline int spmv_grid(int64_t rows, int tpr, int num_sms) {
    int64_t groups_per_block = kBlock / tpr;
    int64_t want = (rows + groups_per_block - 1) / groups_per_block;
    int64_t cap = (int64_t)num_sms * 8;
    if (want < 1) want = 1;
    return (int)(want < cap ? want : cap);
}

// Launch with the matrix's sub-warp width.  x must cover cols (+ halo tail when partitioned).
template <int NS, class RowOp>
int spmv_launch(Context& c, const nsb_mat_s* m, const double* x, RowOp op) {
    CsrView A = view_of(m);
    const int grid = spmv_grid(m->rows, m->threads_per_row, c.num_sms);
    switch (m->threads_per_row) {
        case 2: spmv_csrvec_kernel<2, NS, RowOp><<<grid, kBlock, 0, c.stream>>>(A, x, op, c.partials, c.ticket); break;
        case 4: spmv_csrvec_kernel<4, NS, RowOp><<<grid, kBlock, 0, c.stream>>>(A, x, op, c.partials, c.ticket); break;
        case 8: spmv_csrvec_kernel<8, NS, RowOp><<<grid, kBlock, 0, c.stream>>>(A, x, op, c.partials, c.ticket); break;
        case 16: spmv_csrvec_kernel<16, NS, RowOp><<<grid, kBlock, 0, c.stream>>>(A, x, op, c.partials, c.ticket); break;
        default: spmv_csrvec_kernel<32, NS, RowOp><<<grid, kBlock, 0, c.stream>>>(A, x, op, c.partials, c.ticket); break;
    }
    c.launches++;
    NSB_CUDA(cudaGetLastError());
    return NSB_OK;
}

// y = M x
struct PlainRowOp {
    double* __restrict__ y;
    __device__ bool skip() const { return false; }
    __device__ void row(int64_t r, double ax, double (&)[1]) const { y[r] = ax; }
    __device__ void finish(const double (&)[1]) const {}
};

}  // namespace nsb
