// p2p.cuh -- device side of the peer-memory scalar all-reduce (p2p.cu), as a template so that a solver can run its
// scalar recurrence (Givens rotation, alpha / beta, stop test) in the same one-block kernel, right after the sum.
#pragma once

#include "common.cuh"

namespace nsb {

typedef unsigned long long u64;
constexpr int kArSlots = 8;   // doubles per rank per reduction

__device__ __forceinline__ void st_release_sys(u64* p, u64 v) {
    asm volatile("st.release.sys.global.u64 [%0], %1;" ::"l"(p), "l"(v) : "memory");
}
__device__ __forceinline__ u64 ld_acquire_sys(const u64* p) {
    u64 v;
    asm volatile("ld.acquire.sys.global.u64 %0, [%1];" : "=l"(v) : "l"(p) : "memory");
    return v;
}
__device__ __forceinline__ void st_sys_f64(double* p, double v) {
    asm volatile("st.volatile.global.f64 [%0], %1;" ::"l"(p), "d"(v) : "memory");
}

#ifdef NSB_SYNC_STATS
__device__ __forceinline__ u64 global_ns() {
    u64 t;
    asm volatile("mov.u64 %0, %%globaltimer;" : "=l"(t));
    return t;
}
#define NSB_STATS_T0() const u64 stats_t0__ = global_ns()
#define NSB_STATS_ADD(ptr, slot)                                        \
    do {                                                                \
        if (ptr) {                                                      \
            atomicAdd((ptr) + (slot), global_ns() - stats_t0__);        \
            atomicAdd((ptr) + (slot) + 1, 1ull);                        \
        }                                                               \
    } while (0)
#else
#define NSB_STATS_T0() do { } while (0)
#define NSB_STATS_ADD(ptr, slot) do { } while (0)
#endif

struct P2PReduceView {
    double* const* peer_box;   // per rank: that rank's mailbox
    u64* const* peer_flag;     // per rank: that rank's flag array
    const double* my_box;
    const u64* my_flag;
    u64* seq;
    int me, nranks;
    u64* stats;   // NSB_SYNC_STATS builds; else null
};

// What a SpMV kernel needs to take the halo entries of x straight from this rank's landing buffer (spmv.cuh): the
// kernel visits the chunks that touch no halo column first and polls the neighbours' ready flags only when it
// reaches the others, so the wait for the exchange overlaps the interior rows and no copy kernel is launched.
struct HaloIn {
    const u64* my_ready;   // one word per neighbour, released by that neighbour's k_halo_push
    const u64* seq;        // number of the exchange pushed last on this stream
    const double* land;    // landing area of the vector being multiplied
    int64_t n_owned;       // columns >= n_owned are halo entries, in landing-buffer order
    int nnb;
    const int* order;      // chunk visiting order of the panel, interior chunks first
    int n_interior;
    u64* stats;            // NSB_SYNC_STATS builds; else null
};
struct NoHaloIn {};

struct NoPost {
    __device__ void operator()() const {}
};

// One block.  Every rank stores its `count` partial sums into slot [me] of every peer's mailbox (double-buffered by
// the parity of the sequence number), releases a flag there, polls its own flags and adds the slots in rank order:
// all ranks get the same bits.  `post` then runs in thread 0 with the sums already in vals[].
template <class Post>
__global__ void k_p2p_allreduce(P2PReduceView R, double* vals, int count, Post post) {
    __shared__ double mine[kArSlots];
    const u64 e = *R.seq + 1;
    const int par = (int)(e & 1);
    if (threadIdx.x < count) mine[threadIdx.x] = vals[threadIdx.x];
    __syncthreads();
    const int r = threadIdx.x;
    if (r < R.nranks) {
        double* dst = R.peer_box[r] + ((size_t)par * R.nranks + R.me) * kArSlots;
        for (int k = 0; k < count; ++k) st_sys_f64(dst + k, mine[k]);
        __threadfence_system();
        st_release_sys(R.peer_flag[r] + R.me, e);
    }
    NSB_STATS_T0();
    if (r < R.nranks) {
        while (ld_acquire_sys(R.my_flag + r) < e) {
        }
    }
    __syncthreads();
    if (threadIdx.x == 0) NSB_STATS_ADD(R.stats, 0);
    if (threadIdx.x < count) {
        double s = 0.0;
        for (int q = 0; q < R.nranks; ++q) s += __ldcv(R.my_box + ((size_t)par * R.nranks + q) * kArSlots + threadIdx.x);
        vals[threadIdx.x] = s;
    }
    __syncthreads();
    if (threadIdx.x == 0) {
        *R.seq = e;
        post();
    }
}

P2PReduceView p2p_reduce_view(Context& c);   // p2p.cu
HaloIn p2p_halo_in(const nsb_halo_s* h, int area, const int* order, int n_interior);   // p2p.cu

// Sum `count` (<= kArSlots) doubles over the ranks and run `post` on the result in the same kernel.  Returns false
// (nothing launched) when the context has no peer-memory path: the caller then uses allreduce_sum + its own kernel.
template <class Post>
bool p2p_allreduce_then(Context& c, double* dev, int count, Post post) {
    if (!c.p2p || count > kArSlots) return false;
    const int threads = ((c.nranks > count ? c.nranks : count) + 31) / 32 * 32;
    k_p2p_allreduce<Post><<<1, threads, 0, c.stream>>>(p2p_reduce_view(c), dev, count, post);
    c.launches++;
    return true;
}

}  // namespace nsb
