// cond.cuh -- statically condensed velocity operator of solvers/bramblepasciak_new.py:8-21, 84-103.
//   matA = (I - HT)(A + inner_matrix)(I - H)                          (myAmatrix, :88)
//   harmonic_extension(f) = y + H y + inner_solve r,  r = f + HT f,  y = inverse r   (:12-18)
// Both are chains of SpMVs with one-line row epilogues; the last SpMV of matA hands its row result to the
// caller's row op (SubThenOp), so the dot products the solvers fuse into an `A` apply stay fused.
// Row-partitioned (SURVEY.md 8(e) applied to 8(f) item 1): the five operators hold this rank's rows with columns numbered
// [owned | halo] over ONE shared velocity halo plan (partition.distribute_operators); every SpMV of a chain is preceded by the
// exchange of its input vector (halo_exchange: a no-op for an operator without a plan, so the single-GPU chain is unchanged).
// inner_solve / inner_matrix are element-local: with elements kept inside a slab they reference owned columns only and the
// host side attaches no plan to them.
#pragma once

#include "spmv.cuh"

struct nsb_cond_s {
    nsb::Context* ctx = nullptr;
    nsb_mat_s *A = nullptr, *H = nullptr, *HT = nullptr, *inner_solve = nullptr, *inner_matrix = nullptr;   // not owned
    int64_t n = 0;     // rows (owned rows when row-partitioned)
    int64_t nx = 0;    // length of a vector the chain reads: owned + halo entries (== n on one GPU)
    nsb::DevBuf<double> t, q;   // scratch of the operator apply, nx entries each
};

namespace nsb {

struct SubInOp {   // out = xin - M x
    const double* __restrict__ xin;
    double* __restrict__ out;
    __device__ bool skip() const { return false; }
    __device__ void row(int64_t r, double ax, double (&)[1]) const { out[r] = xin[r] - ax; }
    __device__ void finish(const double (&)[1]) const {}
};

struct AddInOp {   // out = xin + M x   (out may alias xin when xin is not the SpMV input)
    const double* xin;
    double* out;
    __device__ bool skip() const { return false; }
    __device__ void row(int64_t r, double ax, double (&)[1]) const { out[r] = xin[r] + ax; }
    __device__ void finish(const double (&)[1]) const {}
};

struct ScaledAddInOp {   // r = xin + M x ; y = s d .* r     (:12-15 in one pass)
    const double* __restrict__ xin;
    const double* __restrict__ d;
    double s;
    double* __restrict__ r_out;
    double* __restrict__ y_out;
    __device__ bool skip() const { return false; }
    __device__ void row(int64_t r, double ax, double (&)[1]) const {
        const double v = xin[r] + ax;
        r_out[r] = v;
        y_out[r] = s * (d[r] * v);
    }
    __device__ void finish(const double (&)[1]) const {}
};

template <class Inner, int NSA>
struct SubThenOp {   // hands (q - M q)_r to the inner row op
    const double* __restrict__ q;
    Inner in;
    __device__ bool skip() const { return in.skip(); }
    __device__ void row(int64_t r, double ax, double (&acc)[NSA]) const { in.row(r, q[r] - ax, acc); }
    __device__ void finish(const double (&tot)[NSA]) const { in.finish(tot); }
    using BlockFinish = void;   // all threads of the last block: the inner op may end its reduction in a peer all-reduce
    __device__ void finish_block(const double (&tot)[NSA]) const {
        double t[NSA];
#pragma unroll
        for (int k = 0; k < NSA; ++k) t[k] = tot[k];
        finish_last_block(in, t);
    }
};

// fin.row(r, (matA x)_r, acc) for every row.  x is a device vector of K->nx entries (not K->t / K->q) whose first K->n are the
// input; row-partitioned, its halo tail is filled here.
template <int NS, class Inner>
int cond_apply_A(Context& c, nsb_cond_s* K, double* x, Inner fin) {
    NSB_CHECK(halo_exchange(c, K->H->halo, x));
    NSB_CHECK(spmv_launch<0>(c, K->H, x, SubInOp{x, K->t.p}));                        // t = x - H x
    NSB_CHECK(halo_exchange(c, K->A->halo ? K->A->halo : K->inner_matrix->halo, K->t.p));
    NSB_CHECK(spmv_launch<0>(c, K->A, K->t.p, PlainRowOp{K->q.p}));                   // q = A t
    NSB_CHECK(spmv_launch<0>(c, K->inner_matrix, K->t.p, AddInOp{K->q.p, K->q.p}));  // q += inner_matrix t
    NSB_CHECK(halo_exchange(c, K->HT->halo, K->q.p));
    return spmv_launch<NS>(c, K->HT, K->q.p, SubThenOp<Inner, (NS > 0 ? NS : 1)>{K->q.p, fin});
}

// y_out = harmonic_extension(f) with inverse = s * diag(d).  r, y: scratch; none of the five may alias.  Row-partitioned: f, r, y
// hold K->nx entries (their halo tails are filled here), y_out K->n.
inline int cond_apply_pre(Context& c, nsb_cond_s* K, double* f, const double* d, double s, double* r, double* y,
                          double* y_out) {
    NSB_CHECK(halo_exchange(c, K->HT->halo, f));
    NSB_CHECK(spmv_launch<0>(c, K->HT, f, ScaledAddInOp{f, d, s, r, y}));             // r = f + HT f ; y = inverse r
    NSB_CHECK(halo_exchange(c, K->H->halo, y));
    NSB_CHECK(spmv_launch<0>(c, K->H, y, AddInOp{y, y_out}));                         // y_out = y + H y
    NSB_CHECK(halo_exchange(c, K->inner_solve->halo, r));
    return spmv_launch<0>(c, K->inner_solve, r, AddInOp{y_out, y_out});                // y_out += inner_solve r
}

}  // namespace nsb
