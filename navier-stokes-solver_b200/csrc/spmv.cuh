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

#include <stdlib.h>

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
                                                             double* partials, unsigned int* ticket, int chunk_steps) {
    if (op.skip()) return;
    constexpr int NSA = NS > 0 ? NS : 1;
    constexpr int GPB = kBlock / TPR;  // row groups per block
    double acc[NSA];
#pragma unroll
    for (int k = 0; k < NSA; ++k) acc[k] = 0.0;

    const int lane = threadIdx.x & (TPR - 1);
    const int gib = threadIdx.x / TPR;  // group index inside the block
    const int64_t chunk_rows = (int64_t)GPB * chunk_steps;

    // A block owns whole chunks of consecutive rows (chunk, chunk + grid, ...): consecutive rows of an FEM
    // matrix share most of their columns, so the x gathers of a chunk are served from the SM's L1 instead
    // of L2, while the value / index streams bypass L1.  All loop bounds are block-uniform, so the
    // full-mask shuffles below are always executed by all 32 lanes.
    for (int64_t base = (int64_t)blockIdx.x * chunk_rows; base < A.rows; base += (int64_t)gridDim.x * chunk_rows)
    for (int st = 0; st < chunk_steps; ++st) {
        const int64_t r = base + (int64_t)st * GPB + gib;
        if (base + (int64_t)st * GPB >= A.rows) break;
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

inline int spmv_chunk_steps() {
    static int v = -1;
    if (v < 0) {
        const char* e = getenv("NSB_CHUNK_STEPS");
        v = e ? atoi(e) : 16;
        if (v < 1) v = 1;
    }
    return v;
}

inline int spmv_grid(int64_t rows, int tpr, int num_sms) {
    int64_t chunk_rows = (int64_t)(kBlock / tpr) * spmv_chunk_steps();
    int64_t want = (rows + chunk_rows - 1) / chunk_rows;
    int64_t cap = (int64_t)num_sms * 8;
    if (want < 1) want = 1;
    return (int)(want < cap ? want : cap);
}

// ---------------------------------------------------------------------------------------------
// LX kernel: x staged in shared memory per chunk of rows, 16-bit chunk-local column indices (lx.cu).
// Lane l of a TPR-wide group takes nonzeros p0+l, p0+l+TPR, ... so a group reads contiguous runs of the
// value stream (64-bit loads) and of the index stream (16-bit loads) and consecutive columns land
// in consecutive shared-memory banks; four loads per stream are in flight per lane.
struct LxView {
    int R, nchunks;
    const unsigned short* __restrict__ lcol;
    const int* __restrict__ uptr;
    const int* __restrict__ ucol;
};

template <int TPR, int NS, class RowOp>
__global__ void __launch_bounds__(kLxBlock) spmv_lx_kernel(CsrView A, LxView L, const double* __restrict__ x, RowOp op,
                                                           double* partials, unsigned int* ticket) {
    extern __shared__ double sx[];
    if (op.skip()) return;
    constexpr int NSA = NS > 0 ? NS : 1;
    constexpr int GPB = kLxBlock / TPR;
    double acc[NSA];
#pragma unroll
    for (int k = 0; k < NSA; ++k) acc[k] = 0.0;
    const int lane = threadIdx.x & (TPR - 1);
    const int gib = threadIdx.x / TPR;

    for (int chunk = blockIdx.x; chunk < L.nchunks; chunk += gridDim.x) {
        const int u0 = __ldg(L.uptr + chunk);
        const int nu = __ldg(L.uptr + chunk + 1) - u0;
        __syncthreads();  // the previous chunk's gathers are done with sx
        for (int i = threadIdx.x; i < nu; i += 4 * kLxBlock) {
            int c[4];
            double v[4];
#pragma unroll
            for (int k = 0; k < 4; ++k) c[k] = (i + k * kLxBlock < nu) ? __ldg(L.ucol + u0 + i + k * kLxBlock) : 0;
#pragma unroll
            for (int k = 0; k < 4; ++k) v[k] = (i + k * kLxBlock < nu) ? __ldg(x + c[k]) : 0.0;
#pragma unroll
            for (int k = 0; k < 4; ++k)
                if (i + k * kLxBlock < nu) sx[i + k * kLxBlock] = v[k];
        }
        __syncthreads();
        const int64_t r0 = (int64_t)chunk * L.R;
        for (int rb = 0; rb < L.R; rb += GPB) {
            if (r0 + rb >= A.rows) break;  // block-uniform
            const int64_t r = r0 + rb + gib;
            const bool valid = r < A.rows;
            const int p0 = valid ? __ldg(A.rowptr + r) : 0;
            const int p1 = valid ? __ldg(A.rowptr + r + 1) : 0;
            double s = 0.0;
            for (int p = p0 + lane; p < p1; p += 4 * TPR) {
                double v[4];
                unsigned short lc[4];
#pragma unroll
                for (int k = 0; k < 4; ++k) {
                    const bool on = p + k * TPR < p1;
                    v[k] = on ? __ldg(A.values + p + k * TPR) : 0.0;
                    lc[k] = on ? __ldg(L.lcol + p + k * TPR) : (unsigned short)0;
                }
#pragma unroll
                for (int k = 0; k < 4; ++k) s = fma(v[k], sx[lc[k]], s);
            }
#pragma unroll
            for (int o = TPR / 2; o > 0; o >>= 1) s += __shfl_xor_sync(0xffffffffu, s, o);
            if (lane == 0 && valid) op.row(r, s, acc);
        }
    }
    if constexpr (NS > 0) {
        if (grid_sum<NSA>(acc, partials, ticket)) {
            if (threadIdx.x == 0) op.finish(acc);
        }
    }
}

template <int TPR, int NS, class RowOp>
int spmv_lx_launch_t(Context& c, const nsb_mat_s* m, const double* x, RowOp op) {
    static int blocks_per_sm = 0;   // per instantiation
    const size_t smem = sizeof(double) * kLxCap;
    if (blocks_per_sm == 0) {
        NSB_CUDA(cudaFuncSetAttribute(spmv_lx_kernel<TPR, NS, RowOp>, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem));
        int b = 0;
        NSB_CUDA(cudaOccupancyMaxActiveBlocksPerMultiprocessor(&b, spmv_lx_kernel<TPR, NS, RowOp>, kLxBlock, smem));
        blocks_per_sm = b > 0 ? b : 1;
    }
    int grid = c.num_sms * blocks_per_sm;
    if (grid > m->lx_nchunks) grid = m->lx_nchunks;
    if (grid > c.max_grid) grid = c.max_grid;
    LxView L{m->lx_R, m->lx_nchunks, m->lx_lcol.p, m->lx_uptr.p, m->lx_ucol.p};
    spmv_lx_kernel<TPR, NS, RowOp><<<grid, kLxBlock, smem, c.stream>>>(view_of(m), L, x, op, c.partials, c.ticket);
    c.launches++;
    NSB_CUDA(cudaGetLastError());
    return NSB_OK;
}

// Launch with the matrix's sub-warp width.  x must cover cols (+ halo tail when partitioned).
template <int NS, class RowOp>
int spmv_launch(Context& c, const nsb_mat_s* m, const double* x, RowOp op) {
    if (m->lx_R > 0) {
        switch (m->threads_per_row) {
            case 2: return spmv_lx_launch_t<2, NS, RowOp>(c, m, x, op);
            case 4: return spmv_lx_launch_t<4, NS, RowOp>(c, m, x, op);
            case 8: return spmv_lx_launch_t<8, NS, RowOp>(c, m, x, op);
            case 16: return spmv_lx_launch_t<16, NS, RowOp>(c, m, x, op);
            default: return spmv_lx_launch_t<32, NS, RowOp>(c, m, x, op);
        }
    }
    CsrView A = view_of(m);
    const int grid = spmv_grid(m->rows, m->threads_per_row, c.num_sms);
    const int cs = spmv_chunk_steps();
    switch (m->threads_per_row) {
        case 2: spmv_csrvec_kernel<2, NS, RowOp><<<grid, kBlock, 0, c.stream>>>(A, x, op, c.partials, c.ticket, cs); break;
        case 4: spmv_csrvec_kernel<4, NS, RowOp><<<grid, kBlock, 0, c.stream>>>(A, x, op, c.partials, c.ticket, cs); break;
        case 8: spmv_csrvec_kernel<8, NS, RowOp><<<grid, kBlock, 0, c.stream>>>(A, x, op, c.partials, c.ticket, cs); break;
        case 16: spmv_csrvec_kernel<16, NS, RowOp><<<grid, kBlock, 0, c.stream>>>(A, x, op, c.partials, c.ticket, cs); break;
        default: spmv_csrvec_kernel<32, NS, RowOp><<<grid, kBlock, 0, c.stream>>>(A, x, op, c.partials, c.ticket, cs); break;
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
