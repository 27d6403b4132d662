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

#include <type_traits>

#include "common.cuh"
#include "p2p.cuh"

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
    pdl_enter();
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
        if (grid_sum<NSA>(acc, partials, ticket)) finish_last_block(op, acc);
    }
}

// Rows a block walks through per chunk, in steps of GPB rows.  16 steps give the L1 reuse of consecutive FEM rows on
// large operators; small operators (the latency-bound configs of 10^3..10^5 rows) get fewer steps so that the grid
// still covers every SM -- with 16 steps a 5000-row matrix ran on 10 blocks and an SpMV took 15 us instead of 4.
inline int spmv_chunk_steps(int64_t rows, int tpr, int num_sms) {
    static int forced = -2;
    if (forced == -2) {
        const char* e = getenv("NSB_CHUNK_STEPS");
        forced = e ? atoi(e) : -1;
    }
    if (forced > 0) return forced;
    const int64_t gpb = kBlock / tpr;
    int64_t cs = rows / (gpb * (int64_t)num_sms * 2);
    if (cs < 1) cs = 1;
    if (cs > 16) cs = 16;
    return (int)cs;
}

inline int spmv_grid(int64_t rows, int tpr, int num_sms) {
    int64_t chunk_rows = (int64_t)(kBlock / tpr) * spmv_chunk_steps(rows, tpr, num_sms);
    int64_t want = (rows + chunk_rows - 1) / chunk_rows;
    int64_t cap = (int64_t)num_sms * 8;
    if (want < 1) want = 1;
    return (int)(want < cap ? want : cap);
}

// ---------------------------------------------------------------------------------------------
// SELL-LX kernel: sliced-ELL storage sorted inside chunks of R rows, x staged in shared memory per
// chunk, 16-bit chunk-local column indices (format built by sell.cu).
//
//   chunk   = R consecutive rows; its distinct columns are listed once (ucol) and x[ucol] is staged into
//             shared memory with coalesced loads, so the per-nonzero gather costs shared-memory bank
//             cycles instead of L1/L2 sectors and the index stream is 2 B per nonzero;
//   slice   = 32/T rows of a chunk after sorting the chunk's rows by length; T lanes share a row.  A slice
//             is stored as pair-blocks of 32 x (2 values) so that every lane issues 128-bit value loads and
//             32-bit index loads and a warp reads 512 B / 128 B contiguous per load;
//   warp    = takes the slices of a chunk in snake order (longest with shortest) so the warps of a block
//             finish together; no row pointers are read.
// Row sums land in shared memory (sy) in natural row order; after a barrier the RowOp epilogue runs
// with thread t on row r0 + t, so every fused vector read/write is coalesced.
constexpr int kSellBlock = 256;
constexpr int kSellWarps = kSellBlock / 32;
constexpr int kSellMaxSmem = 160 * 1024;
#ifndef NSB_SELL_UNROLL
#define NSB_SELL_UNROLL 8
#endif
#ifndef NSB_SELL_MINBLOCKS
#define NSB_SELL_MINBLOCKS 4
#endif

struct SellView {
    int64_t rows;   // rows of the panel
    int64_t row0;   // first operator row of the panel
    int R, log2R, nchunks, nsl;   // rows per chunk, chunks, slices per chunk (R * T / 32)
    int cap;                      // staged x entries a chunk may need (max distinct columns, padded)
    int sell_flags;               // Context::sell_flags
    const double* __restrict__ vals;
    const unsigned short* __restrict__ lcol;
    const unsigned int* __restrict__ off;      // per slice, in pair-blocks (64 elements)
    const unsigned short* __restrict__ perm;   // sorted position -> row inside the chunk
    const int* __restrict__ uptr;
    const int* __restrict__ ucol;
    // runs of consecutive distinct columns (rptr == nullptr: use the ucol list)
    const int* __restrict__ rptr;
    const int* __restrict__ rstart;
    const int* __restrict__ rpos;
};

__device__ __forceinline__ unsigned int ld_stream_b32(const void* p) {
    unsigned int r;
    asm volatile("ld.global.nc.L1::no_allocate.b32 %0, [%1];" : "=r"(r) : "l"(p));
    return r;
}

// 8-byte asynchronous copy global -> shared (LDGSTS through L1, like the __ldg it replaces) and the wait for all copies this
// thread issued; the block barrier after the wait publishes them to the other threads
__device__ __forceinline__ void cp_async_f64(double* smem_dst, const double* gmem_src) {
    const unsigned int d = (unsigned int)__cvta_generic_to_shared(smem_dst);
    asm volatile("cp.async.ca.shared.global [%0], [%1], 8;" ::"r"(d), "l"(gmem_src) : "memory");
}
__device__ __forceinline__ void cp_async_b32(unsigned int* smem_dst, const unsigned int* gmem_src) {
    const unsigned int d = (unsigned int)__cvta_generic_to_shared(smem_dst);
    asm volatile("cp.async.ca.shared.global [%0], [%1], 4;" ::"r"(d), "l"(gmem_src) : "memory");
}
__device__ __forceinline__ void cp_async_wait_all() { asm volatile("cp.async.wait_all;" ::: "memory"); }

// x[c] for the staging phase.  Row-partitioned operators with the in-kernel wait (HaloIn) take the halo columns
// from the landing buffer the neighbours store into (coherent loads: the data arrived while this kernel ran).
template <class HI>
__device__ __forceinline__ double ld_x(const double* __restrict__ x, const HI& H, int c) {
    if constexpr (!std::is_same<HI, NoHaloIn>::value) {
        if (c >= H.n_owned) return __ldcv(H.land + (c - H.n_owned));
    }
    return __ldg(x + c);
}

// Optional parts of a RowOp the SELL kernel uses when present:
//   Pre / load(i) / apply(i, ax, pre, acc)   row(i, ax, acc) == apply(i, ax, load(i), acc), split so that the epilogue
//                          can have the operand loads of two rows per thread in flight before the first use (the
//                          epilogue is latency-bound: +1.5 % BPCG iterations/s at 20M dofs, +2 % at 2.5M; touching the
//                          operand lines with prefetch.global.L2 while x is staged was measured too and LOSES 1.3 %)
template <class Op, class = void>
struct has_load : std::false_type {};
template <class Op>
struct has_load<Op, std::void_t<typename Op::Pre>> : std::true_type {};
template <class Op, class = void>
struct pre_of {
    using type = int;
};
template <class Op>
struct pre_of<Op, std::void_t<typename Op::Pre>> {
    using type = typename Op::Pre;
};

// finish_block(tot): like finish(tot) (totals valid in thread 0) but called by ALL threads of the last block, for row ops whose
// reduction ends in a peer all-reduce run by that block (p2p.cuh: reduce_finish)
template <class Op, class = void>
struct has_finish_block : std::false_type {};
template <class Op>
struct has_finish_block<Op, std::void_t<typename Op::BlockFinish>> : std::true_type {};   // marker: `using BlockFinish = void;`
// push_out(): the row op ships interface rows of a vector it writes to the neighbours (p2p.cuh: PushOut); the kernel runs the
// protocol around the row epilogues (push_begin / push_wait_consumed / push_end), the op calls push_store for its rows
template <class Op, class = void>
struct has_push : std::false_type {};
template <class Op>
struct has_push<Op, std::void_t<decltype(std::declval<const Op&>().push_out())>> : std::true_type {};

template <class Op, int NSA>
__device__ __forceinline__ void finish_last_block(const Op& op, double (&acc)[NSA]) {
    if constexpr (has_finish_block<Op>::value) op.finish_block(acc);
    else {
        if (threadIdx.x == 0) op.finish(acc);
    }
}

template <int T, int NS, class RowOp, class HI>
__global__ void __launch_bounds__(kSellBlock, NSB_SELL_MINBLOCKS) spmv_sell_kernel(SellView S, const double* __restrict__ x, RowOp op,
                                                               double* partials, unsigned int* ticket, HI H) {
    extern __shared__ double smem[];
    constexpr bool HALO = !std::is_same<HI, NoHaloIn>::value;
    if (S.sell_flags & 1) pdl_trigger();
    // Before the dependency wait: pull the structure and the head of the value / index streams of this block's first
    // chunk towards L2.  None of it is ever written after the operator was built, so it may be fetched while the
    // previous kernel of the stream is still draining; x, the row-op operands and the reduction scratch come after.
    if ((S.sell_flags & 2) && blockIdx.x < S.nchunks) {
        int chunk0 = blockIdx.x;
        if constexpr (HALO) chunk0 = __ldg(H.order + blockIdx.x);
        const unsigned int* offp = S.off + (size_t)chunk0 * S.nsl;
        const unsigned int b0 = __ldg(offp), b1 = __ldg(offp + S.nsl);   // pair-blocks: 512 B of values, 128 B of indices each
        const size_t line = threadIdx.x;                                 // one 128-byte line per thread
        if (line * 8 < (size_t)(b1 - b0) * 32) prefetch_l2(S.vals + (size_t)b0 * 64 + line * 16);
        if (line < 64 && line * 32 < (size_t)(b1 - b0) * 32) prefetch_l2(S.lcol + (size_t)b0 * 64 + line * 64);
        if (line * 32 <= (size_t)S.nsl) prefetch_l2(offp + line * 32);
        if (line * 64 < (size_t)S.R) prefetch_l2(S.perm + ((size_t)chunk0 << S.log2R) + line * 64);
        if (threadIdx.x == 0) prefetch_l2(S.uptr + chunk0);
        if (S.rptr != nullptr) {
            const int q0 = __ldg(S.rptr + chunk0), q1 = __ldg(S.rptr + chunk0 + 1);
            if ((int)line * 32 < q1 - q0) {
                prefetch_l2(S.rstart + q0 + line * 32);
                prefetch_l2(S.rpos + q0 + line * 32);
            }
        } else {
            const int q0 = __ldg(S.uptr + chunk0), q1 = __ldg(S.uptr + chunk0 + 1);
            for (int i = (int)line * 32; i < q1 - q0; i += kSellBlock * 32) prefetch_l2(S.ucol + q0 + i);
        }
    }
    pdl_wait();
    if (op.skip()) return;
    u64 exch = 0;
    bool landed = false;
    if constexpr (HALO) exch = *H.seq;   // written by the push of this exchange, earlier in the stream
    u64 push_e = 0;
    bool push_open = false;   // the neighbours have consumed the previous exchange: this block may store into their landing buffers
    if constexpr (has_push<RowOp>::value) {
        if (op.push_out().nruns > 0) push_e = push_begin(op.push_out());
    }
    double* sx = smem;
    double* sy = smem + S.cap;
    unsigned int* soff = reinterpret_cast<unsigned int*>(sy + S.R);
    constexpr int NSA = NS > 0 ? NS : 1;
    constexpr int RPS = 32 / T;   // rows per slice
    constexpr int U = NSB_SELL_UNROLL;
    double acc[NSA];
#pragma unroll
    for (int k = 0; k < NSA; ++k) acc[k] = 0.0;
    const int lane = threadIdx.x & 31, warp = threadIdx.x >> 5;

    // Asynchronous staging of one chunk (slice offsets and x entries): every copy global -> shared is a cp.async (LDGSTS)
    // that occupies no register and nothing waits on, so a warp issues the copies of ALL its runs back to back and the
    // chunk pays one memory latency for its x entries instead of one per pair of runs.
    auto stage_async = [&](int ch) {
        const int u0 = __ldg(S.uptr + ch);
        for (int i = threadIdx.x; i <= S.nsl; i += kSellBlock) cp_async_b32(soff + i, S.off + (size_t)ch * S.nsl + i);
        if (S.rptr != nullptr) {
            const int r0 = __ldg(S.rptr + ch), r1 = __ldg(S.rptr + ch + 1);
            for (int rb = r0 + warp; rb < r1; rb += 32 * kSellWarps) {
                const int mine = rb + lane * kSellWarps;
                const int cs_l = mine < r1 ? __ldg(S.rstart + mine) : 0;
                const int p0_l = mine < r1 ? __ldg(S.rpos + mine) : 0;
                const int p1_l = mine < r1 ? __ldg(S.rpos + mine + 1) : 0;
                const int cnt = min(32, (r1 - rb + kSellWarps - 1) / kSellWarps);
                for (int k = 0; k < cnt; ++k) {
                    const int cs = __shfl_sync(0xffffffffu, cs_l, k);
                    const int p0 = __shfl_sync(0xffffffffu, p0_l, k) - u0;
                    const int len = __shfl_sync(0xffffffffu, p1_l, k) - u0 - p0;
                    if constexpr (HALO) {
                        // halo columns come from the landing buffer the neighbours store into: coherent loads, no L1.  Owned and
                        // halo columns are numbered apart, so a run is one or the other (a run across the border is cut here)
                        if (cs + len <= H.n_owned) {          // (warp-uniform) an owned run: the same loop as the plain kernel
                            for (int l = lane; l < len; l += 32) cp_async_f64(sx + p0 + l, x + cs + l);
                        } else if (cs >= H.n_owned) {
                            for (int l = lane; l < len; l += 32) sx[p0 + l] = __ldcv(H.land + (cs + l - H.n_owned));
                        } else {                              // a run across the owned | halo border
                            for (int l = lane; l < len; l += 32) {
                                if (cs + l >= H.n_owned) sx[p0 + l] = __ldcv(H.land + (cs + l - H.n_owned));
                                else cp_async_f64(sx + p0 + l, x + cs + l);
                            }
                        }
                    } else {
                        for (int l = lane; l < len; l += 32) cp_async_f64(sx + p0 + l, x + cs + l);
                    }
                }
            }
        } else {
            const int nu = __ldg(S.uptr + ch + 1) - u0;
            for (int i = threadIdx.x; i < nu; i += 4 * kSellBlock) {
                int c[4];
#pragma unroll
                for (int k = 0; k < 4; ++k) c[k] = (i + k * kSellBlock < nu) ? __ldg(S.ucol + u0 + i + k * kSellBlock) : 0;
#pragma unroll
                for (int k = 0; k < 4; ++k)
                    if (i + k * kSellBlock < nu) {
                        if constexpr (HALO) {
                            if (c[k] >= H.n_owned) {
                                sx[i + k * kSellBlock] = __ldcv(H.land + (c[k] - H.n_owned));
                                continue;
                            }
                        }
                        cp_async_f64(sx + i + k * kSellBlock, x + c[k]);
                    }
            }
        }
    };
    const bool async_stage = (S.sell_flags & 8) != 0;
    // with bit 4 the next chunk of the block is staged right behind the compute phase, so its latency runs under the
    // (latency-bound) row epilogue of the current chunk; one buffer is enough: the compute phase is done with sx
    // row-partitioned: the wait for the halo sits at the loop top, so a chunk with halo columns is staged ahead only once the
    // neighbours' entries have landed; interior chunks (the bulk) are always staged ahead
    const bool pipelined = async_stage && (S.sell_flags & 16);
    bool staged = false;   // this chunk's copies were issued during the previous chunk's epilogue

    // row-partitioned: the chunk visiting order (interior chunks first) is read ONE ITERATION AHEAD -- a load of order[ci] at the
    // top of the iteration sat in front of every other load of the chunk (uptr, the runs, the streams: +6..19 % per SpMV)
    int chunk_ahead = 0;
    if constexpr (HALO) chunk_ahead = blockIdx.x < S.nchunks ? __ldg(H.order + blockIdx.x) : 0;
    for (int ci = blockIdx.x; ci < S.nchunks; ci += gridDim.x) {
        int chunk = ci;
        if constexpr (HALO) {
            chunk = chunk_ahead;
            chunk_ahead = ci + (int)gridDim.x < S.nchunks ? __ldg(H.order + ci + (int)gridDim.x) : 0;
            if (ci >= H.n_interior && !landed) {
                // first chunk with halo columns: the neighbours' entries of this exchange must have landed.  The
                // barrier below orders the other threads of the block behind these acquires.
                if (threadIdx.x < H.nnb) {
                    NSB_STATS_T0();
                    spin_until(H.my_ready + threadIdx.x, exch, H.guard, kFaultHaloReady);
                    if (blockIdx.x == 0 && threadIdx.x == 0) NSB_STATS_ADD(H.stats, 4);
                }
                landed = true;
            }
        }
        if (staged) {
            // nothing to issue
        } else if (async_stage) {
            __syncthreads();   // previous chunk: gathers done with sx, epilogue done with sy
            stage_async(chunk);
        } else {
        const int u0 = __ldg(S.uptr + chunk);
        const int nu = __ldg(S.uptr + chunk + 1) - u0;
        __syncthreads();   // previous chunk: gathers done with sx, epilogue done with sy
        for (int i = threadIdx.x; i <= S.nsl; i += kSellBlock) soff[i] = __ldg(S.off + (size_t)chunk * S.nsl + i);
        if (S.rptr != nullptr) {
            // contiguous copies: lane k of a warp holds the descriptor of the warp's k-th run, runs are then
            // copied two at a time with up to 8 coalesced loads in flight per lane
            const int r0 = __ldg(S.rptr + chunk), r1 = __ldg(S.rptr + chunk + 1);
            for (int rb = r0 + warp; rb < r1; rb += 32 * kSellWarps) {
                const int mine = rb + lane * kSellWarps;
                const int cs_l = mine < r1 ? __ldg(S.rstart + mine) : 0;
                const int p0_l = mine < r1 ? __ldg(S.rpos + mine) : 0;
                const int p1_l = mine < r1 ? __ldg(S.rpos + mine + 1) : 0;
                const int cnt = min(32, (r1 - rb + kSellWarps - 1) / kSellWarps);
                for (int k = 0; k < cnt; k += 2) {
                    int cs[2], p0[2], len[2];
#pragma unroll
                    for (int q = 0; q < 2; ++q) {
                        const int kk = min(k + q, 31);
                        cs[q] = __shfl_sync(0xffffffffu, cs_l, kk);
                        p0[q] = __shfl_sync(0xffffffffu, p0_l, kk) - u0;
                        len[q] = (k + q < cnt) ? __shfl_sync(0xffffffffu, p1_l, kk) - u0 - p0[q] : 0;
                    }
                    const int longest = max(len[0], len[1]);
                    for (int l = lane; l < longest; l += 128) {
                        double v[2][4];
#pragma unroll
                        for (int q = 0; q < 2; ++q)
#pragma unroll
                            for (int j = 0; j < 4; ++j) v[q][j] = (l + 32 * j < len[q]) ? ld_x(x, H, cs[q] + l + 32 * j) : 0.0;
#pragma unroll
                        for (int q = 0; q < 2; ++q)
#pragma unroll
                            for (int j = 0; j < 4; ++j)
                                if (l + 32 * j < len[q]) sx[p0[q] + l + 32 * j] = v[q][j];
                    }
                }
            }
        } else
        for (int i = threadIdx.x; i < nu; i += 4 * kSellBlock) {
            int c[4];
            double v[4];
#pragma unroll
            for (int k = 0; k < 4; ++k) c[k] = (i + k * kSellBlock < nu) ? __ldg(S.ucol + u0 + i + k * kSellBlock) : 0;
#pragma unroll
            for (int k = 0; k < 4; ++k) v[k] = (i + k * kSellBlock < nu) ? ld_x(x, H, c[k]) : 0.0;
#pragma unroll
            for (int k = 0; k < 4; ++k)
                if (i + k * kSellBlock < nu) sx[i + k * kSellBlock] = v[k];
        }
        }
        if (async_stage) cp_async_wait_all();
        __syncthreads();   // x and the slice offsets of this chunk are in shared memory; the previous epilogue is done with sy
        for (int base = 0, pass = 0; base < S.nsl; base += kSellWarps, ++pass) {
            // snake order over the length-sorted slices
            const int cnt = min(kSellWarps, S.nsl - base);
            if (warp >= cnt) continue;
            const int sl = base + ((pass & 1) ? cnt - 1 - warp : warp);
            const unsigned int o0 = soff[sl], o1 = soff[sl + 1];
            const int k = sl * RPS + lane / T;   // sorted position inside the chunk
            const unsigned short lr = __ldg(S.perm + ((size_t)chunk << S.log2R) + k);
            const double2* vp = reinterpret_cast<const double2*>(S.vals) + (size_t)o0 * 32 + lane;
            const unsigned int* cp = reinterpret_cast<const unsigned int*>(S.lcol) + (size_t)o0 * 32 + lane;
            const int n = (int)(o1 - o0);
            double s0 = 0.0, s1 = 0.0;
            // U predicated pair-loads (128-bit values + 32-bit indices) in flight per lane
            for (int j = 0; j < n; j += U) {
                double2 v[U];
                unsigned int c[U];
#pragma unroll
                for (int k2 = 0; k2 < U; ++k2) {
                    const bool on = j + k2 < n;
                    v[k2] = on ? ld_stream_f64x2(reinterpret_cast<const double*>(vp + (size_t)(j + k2) * 32)) : make_double2(0.0, 0.0);
                    c[k2] = on ? ld_stream_b32(cp + (size_t)(j + k2) * 32) : 0u;
                }
#pragma unroll
                for (int k2 = 0; k2 < U; ++k2) {
                    s0 = fma(v[k2].x, sx[c[k2] & 0xffffu], s0);
                    s1 = fma(v[k2].y, sx[c[k2] >> 16], s1);
                }
            }
            double s = s0 + s1;
#pragma unroll
            for (int o = T / 2; o > 0; o >>= 1) s += __shfl_xor_sync(0xffffffffu, s, o);
            if ((lane & (T - 1)) == 0) sy[lr] = s;
        }
        __syncthreads();
        staged = false;
        if (pipelined && ci + (int)gridDim.x < S.nchunks) {
            const int cn = ci + (int)gridDim.x;
            if constexpr (HALO) {
                if (cn < H.n_interior || landed) {
                    stage_async(chunk_ahead);
                    staged = true;
                }
            } else {
                stage_async(cn);
                staged = true;
            }
        }
        const int64_t r0 = (int64_t)chunk << S.log2R;
        if constexpr (has_push<RowOp>::value) {
            // a chunk with rows the op ships to a neighbour (one flag byte per 256 rows): its epilogue also stores those rows into the
            // neighbours' landing buffers; before the first such chunk of this block, wait (bounded) until they consumed the previous exchange
            if (op.push_out().nruns > 0 && push_touches(op.push_out(), S.row0 + r0, S.row0 + min(r0 + (int64_t)S.R, S.rows))) {   // (the last chunk is padded)
                if (!push_open) {
                    push_wait_consumed(op.push_out(), push_e);
                    push_open = true;
                }
                for (int t = threadIdx.x; t < S.R; t += kSellBlock) {
                    const int64_t r = r0 + t;
                    if (r < S.rows) op.row_pushing(S.row0 + r, sy[t], acc);
                }
                continue;
            }
        }
        if constexpr (has_load<RowOp>::value) {
            if (S.sell_flags & 4) {
                using Pre = typename pre_of<RowOp>::type;
                for (int t = threadIdx.x; t < S.R; t += 2 * kSellBlock) {
                    const int64_t ra = r0 + t, rb = ra + kSellBlock;
                    const bool va = ra < S.rows, vb = t + kSellBlock < S.R && rb < S.rows;
                    Pre pa{}, pb{};
                    if (va) pa = op.load(S.row0 + ra);
                    if (vb) pb = op.load(S.row0 + rb);
                    if (va) op.apply(S.row0 + ra, sy[t], pa, acc);
                    if (vb) op.apply(S.row0 + rb, sy[t + kSellBlock], pb, acc);
                }
                continue;
            }
        }
        for (int t = threadIdx.x; t < S.R; t += kSellBlock) {
            const int64_t r = r0 + t;
            if (r < S.rows) op.row(S.row0 + r, sy[t], acc);
        }
    }
    if constexpr (NS > 0) {
        if (grid_sum<NSA>(acc, partials, ticket)) finish_last_block(op, acc);
    }
    if constexpr (has_push<RowOp>::value) {
        if (op.push_out().nruns > 0) push_end(op.push_out(), push_e);
    }
}

inline SellView sell_view_of(const SellData* m) {
    return SellView{m->nrows, m->row0, m->sell_R, m->sell_log2R, m->sell_nchunks, m->sell_R * m->sell_T / 32, m->sell_cap, 0,
                    m->sell_vals.p, m->sell_lcol.p, m->sell_off.p, m->sell_perm.p, m->sell_uptr.p, m->sell_ucol.p,
                    m->sell_nruns > 0 ? m->sell_rptr.p : nullptr, m->sell_rstart.p, m->sell_rpos.p};
}

inline size_t sell_smem_bytes(const SellData* m) {
    const int nsl = m->sell_R * m->sell_T / 32;
    return sizeof(double) * ((size_t)m->sell_cap + m->sell_R) + sizeof(unsigned int) * ((size_t)nsl + 2);
}

template <int T, int NS, class RowOp, class HI>
int spmv_sell_launch_t(Context& c, const SellData* m, const double* x, RowOp op, HI H) {
    // blocks per SM of this instantiation at this shared-memory size, remembered PER CONTEXT (a context belongs to one device and
    // is driven by one host thread at a time); the first use on a context also raises the kernel's shared-memory limit there
    const size_t smem = sell_smem_bytes(m);
    const void* kern = (const void*)spmv_sell_kernel<T, NS, RowOp, HI>;
    int bps = 0;
    bool seen_kernel = false;
    for (const auto& e : c.sell_occupancy) {
        if (e.first.first != kern) continue;
        seen_kernel = true;
        if (e.first.second == smem) {
            bps = e.second;
            break;
        }
    }
    if (bps == 0) {
        if (!seen_kernel)
            NSB_CUDA(cudaFuncSetAttribute(spmv_sell_kernel<T, NS, RowOp, HI>, cudaFuncAttributeMaxDynamicSharedMemorySize, kSellMaxSmem));
        int b = 0;
        NSB_CUDA(cudaOccupancyMaxActiveBlocksPerMultiprocessor(&b, spmv_sell_kernel<T, NS, RowOp, HI>, kSellBlock, smem));
        bps = b > 0 ? b : 1;
        c.sell_occupancy.push_back({{kern, smem}, bps});
    }
    int grid = c.num_sms * bps;
    if (grid > m->sell_nchunks) grid = m->sell_nchunks;
    if (grid > c.max_grid) grid = c.max_grid;
    SellView V = sell_view_of(m);
    V.sell_flags = c.sell_flags;
    NSB_CUDA(launch_k(c, 1, spmv_sell_kernel<T, NS, RowOp, HI>, grid, kSellBlock, smem, V, x, op, c.partials, c.ticket, H));
    c.launches++;
    return NSB_OK;
}

template <int NS, class RowOp, class HI = NoHaloIn>
int spmv_sell_panel(Context& c, const SellData* m, const double* x, RowOp op, HI H = HI{}) {
    switch (m->sell_T) {
        case 1: return spmv_sell_launch_t<1, NS, RowOp, HI>(c, m, x, op, H);
        case 2: return spmv_sell_launch_t<2, NS, RowOp, HI>(c, m, x, op, H);
        case 4: return spmv_sell_launch_t<4, NS, RowOp, HI>(c, m, x, op, H);
        case 8: return spmv_sell_launch_t<8, NS, RowOp, HI>(c, m, x, op, H);
        case 16: return spmv_sell_launch_t<16, NS, RowOp, HI>(c, m, x, op, H);
        default: return spmv_sell_launch_t<32, NS, RowOp, HI>(c, m, x, op, H);
    }
}

// Row panels run as consecutive launches.  The fused reductions of a multi-panel operator are chained: every
// launch but the last parks its totals in c.chain, the last one adds them (fixed order) and runs the inner finish.
template <class Inner, int NSA>
struct ChainOp {
    Inner in;
    double* chain;
    int first, last;
    __device__ bool skip() const { return in.skip(); }
    __device__ void row(int64_t r, double ax, double (&acc)[NSA]) const { in.row(r, ax, acc); }
    template <class I = Inner>
    __device__ typename pre_of<I>::type load(int64_t r) const { return in.load(r); }
    template <class P>
    __device__ void apply(int64_t r, double ax, const P& p, double (&acc)[NSA]) const { in.apply(r, ax, p, acc); }
    __device__ void finish(const double (&tot)[NSA]) const {
        double t[NSA];
#pragma unroll
        for (int k = 0; k < NSA; ++k) t[k] = first ? tot[k] : chain[k] + tot[k];
        if (last) in.finish(t);
        else {
#pragma unroll
            for (int k = 0; k < NSA; ++k) chain[k] = t[k];
        }
    }
    // all threads of the last block of this panel's launch; the totals are valid in thread 0 only
    using BlockFinish = void;
    __device__ void finish_block(const double (&tot)[NSA]) const {
        double t[NSA];
#pragma unroll
        for (int k = 0; k < NSA; ++k) t[k] = first ? tot[k] : chain[k] + tot[k];
        if (last) finish_last_block(in, t);
        else if (threadIdx.x == 0) {
#pragma unroll
            for (int k = 0; k < NSA; ++k) chain[k] = t[k];
        }
    }
};

template <class Inner, int NSA>
struct has_push<ChainOp<Inner, NSA>, void> : std::false_type {};   // a multi-panel SpMV never pushes (one push per exchange)
template <class Inner, int NSA>
struct has_load<ChainOp<Inner, NSA>, void> : has_load<Inner> {};
template <class Inner, int NSA>
struct pre_of<ChainOp<Inner, NSA>, void> : pre_of<Inner> {};

// Launch with the matrix's storage.  x must cover cols (+ halo tail when partitioned).
template <int NS, class RowOp>
int spmv_launch(Context& c, const nsb_mat_s* m, const double* x, RowOp op) {
    const size_t np = m->panels.size();
    if (np == 1) return spmv_sell_panel<NS, RowOp>(c, m->panels[0].get(), x, op);
    if (np > 1) {
        for (size_t k = 0; k < np; ++k) {
            if constexpr (NS > 0) {
                NSB_CHECK((spmv_sell_panel<NS, ChainOp<RowOp, NS>>(c, m->panels[k].get(), x,
                                                                   ChainOp<RowOp, NS>{op, c.chain, k == 0, k + 1 == np})));
            } else {
                NSB_CHECK((spmv_sell_panel<NS, RowOp>(c, m->panels[k].get(), x, op)));
            }
        }
        return NSB_OK;
    }
    CsrView A = view_of(m);
    const int grid = spmv_grid(m->rows, m->threads_per_row, c.num_sms);
    const int cs = spmv_chunk_steps(m->rows, m->threads_per_row, c.num_sms);
    switch (m->threads_per_row) {
        case 2: NSB_CUDA(launch_k(c, 1, spmv_csrvec_kernel<2, NS, RowOp>, grid, kBlock, 0, A, x, op, c.partials, c.ticket, cs)); break;
        case 4: NSB_CUDA(launch_k(c, 1, spmv_csrvec_kernel<4, NS, RowOp>, grid, kBlock, 0, A, x, op, c.partials, c.ticket, cs)); break;
        case 8: NSB_CUDA(launch_k(c, 1, spmv_csrvec_kernel<8, NS, RowOp>, grid, kBlock, 0, A, x, op, c.partials, c.ticket, cs)); break;
        case 16: NSB_CUDA(launch_k(c, 1, spmv_csrvec_kernel<16, NS, RowOp>, grid, kBlock, 0, A, x, op, c.partials, c.ticket, cs)); break;
        default: NSB_CUDA(launch_k(c, 1, spmv_csrvec_kernel<32, NS, RowOp>, grid, kBlock, 0, A, x, op, c.partials, c.ticket, cs)); break;
    }
    c.launches++;
    return NSB_OK;
}

// SpMV of a row-partitioned SELL operator whose halo entries were PUSHED (p2p_halo_push) but not yet waited for: the
// kernel takes them from landing area `area` and waits for the neighbours inside (requires halo_fused_ok(c, m)).
template <int NS, class RowOp>
int spmv_launch_halo_in(Context& c, const nsb_mat_s* m, const double* x, RowOp op, int area) {
    const size_t np = m->panels.size();
    for (size_t k = 0; k < np; ++k) {
        const SellData* P = m->panels[k].get();
        const HaloIn H = p2p_halo_in(m->halo, area, P->halo_order.p, P->halo_n_interior);
        if constexpr (NS > 0) {
            if (np > 1) {
                NSB_CHECK((spmv_sell_panel<NS, ChainOp<RowOp, NS>, HaloIn>(c, P, x, ChainOp<RowOp, NS>{op, c.chain, k == 0, k + 1 == np}, H)));
                continue;
            }
        }
        NSB_CHECK((spmv_sell_panel<NS, RowOp, HaloIn>(c, P, x, op, H)));
    }
    return NSB_OK;
}

// The operator's halo exchange followed by its SpMV.  With the peer-memory path and SELL storage: push + SpMV with
// the wait inside the kernel; otherwise exchange (push + wait/copy kernels, or NCCL) + SpMV.
template <int NS, class RowOp>
int spmv_exchange_launch(Context& c, const nsb_mat_s* m, double* x, RowOp op) {
    if (halo_fused_ok(c, m)) {
        NSB_CHECK(p2p_halo_push(c, m->halo, x));
        return spmv_launch_halo_in<NS>(c, m, x, op, 0);
    }
    NSB_CHECK(halo_exchange(c, m->halo, x));
    return spmv_launch<NS>(c, m, x, op);
}

// y = M x
struct PlainRowOp {
    double* __restrict__ y;
    __device__ bool skip() const { return false; }
    __device__ void row(int64_t r, double ax, double (&)[1]) const { y[r] = ax; }
    __device__ void finish(const double (&)[1]) const {}
};

}  // namespace nsb
