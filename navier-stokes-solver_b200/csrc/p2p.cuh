// p2p.cuh -- device side of the peer-memory exchanges (p2p.cu): the scalar all-reduce as a block-level function, so that the
// LAST BLOCK of the kernel that produced the partial sums runs it together with the solver's scalar recurrence (Givens rotation,
// alpha / beta, stop test) -- no kernel of its own; and the halo push as row-level stores, so that the kernel that PRODUCES a
// vector ships its interface entries from its own epilogue -- no push kernel either.  Every wait on a peer flag is bounded: after
// Context::spin_timeout_ns a fault word is set, the wait is abandoned and the call returns NSB_ERR_NCCL (a dead rank must not
// hang the others inside a kernel).
#pragma once

#include "common.cuh"

namespace nsb {

typedef unsigned long long u64;
constexpr int kArSlots = 8;     // doubles per rank per reduction
constexpr int kPushRuns = 16;   // runs of consecutive owned rows a fused push can describe (x-slabs: one per component and side)

// fault codes written to Context::fault by a wait that timed out
constexpr int kFaultAllreduce = 1, kFaultHaloConsumed = 2, kFaultHaloReady = 3;

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
__device__ __forceinline__ u64 global_ns() {
    u64 t;
    asm volatile("mov.u64 %0, %%globaltimer;" : "=l"(t));
    return t;
}

// what a bounded wait needs: the fault word of the context and the bound
struct SpinGuard {
    int* fault;        // device word; nonzero: an exchange timed out somewhere on this rank
    u64 timeout_ns;    // 0: wait forever
};

// Poll *flag until it reaches `want`.  Returns false (and raises the fault word) when the bound is exceeded or a fault is
// already up -- the caller goes on with whatever data it has; the host turns the fault into an error code.
__device__ __forceinline__ bool spin_until(const u64* flag, u64 want, const SpinGuard& g, int code) {
    if (ld_acquire_sys(flag) >= want) return true;
    const u64 t0 = global_ns();
    unsigned int polls = 0;
    while (ld_acquire_sys(flag) < want) {
        if ((++polls & 63u) == 0 && g.timeout_ns) {
            if (*(volatile int*)g.fault != 0) return false;
            if (global_ns() - t0 > g.timeout_ns) {
                atomicCAS(g.fault, 0, code);
                return false;
            }
        }
    }
    return true;
}

#ifdef NSB_SYNC_STATS
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
    SpinGuard guard;
};

// How the total of a fused reduction is finished (one value per solver call, carried by the row ops / vector kernels):
//   kRedLocal  single GPU: the last block runs the scalar step on its own totals
//   kRedPeer   row-partitioned, peer memory: the last block exchanges the partial sums with every rank through the mailboxes
//              (p2p_allreduce_block) and then runs the scalar step -- every rank computes the same bits
//   kRedStore  row-partitioned, NCCL fallback: the partial sums are left in the state; ncclAllReduce + a scalar kernel follow
enum RedKind { kRedLocal = 0, kRedPeer = 1, kRedStore = 2 };
struct RedMode {
    int kind;
    P2PReduceView R;
};

// What a SpMV kernel needs to take the halo entries of x straight from this rank's landing buffer (spmv.cuh): the
// kernel visits the chunks that touch no halo column first and polls the neighbours' ready flags only when it
// reaches the others, so the wait for the exchange overlaps the interior rows and no copy kernel is launched.
struct HaloIn {
    const u64* my_ready;   // one word per neighbour, released by that neighbour's push
    const u64* seq;        // number of the exchange pushed last on this stream
    const double* land;    // landing area of the vector being multiplied
    int64_t n_owned;       // columns >= n_owned are halo entries, in landing-buffer order
    int nnb;
    const int* order;      // chunk visiting order of the panel, interior chunks first
    int n_interior;
    u64* stats;            // NSB_SYNC_STATS builds; else null
    SpinGuard guard;
};
struct NoHaloIn {};

struct NoPost {
    __device__ void operator()() const {}
};

// Called by ALL threads of one block (blockDim.x >= nranks, >= count).  vals[0..count) in global memory hold this rank's partial
// sums (written before the call, made visible to the block by the barrier inside); on return they hold the totals.  Every rank
// stores its sums into slot [me] of every peer's mailbox (double-buffered by the parity of the sequence number), releases a flag
// there, polls its own flags and adds the slots in rank order: all ranks get the same bits.
__device__ __forceinline__ void p2p_allreduce_block(const P2PReduceView& R, double* vals, int count) {
    __shared__ double mine[kArSlots];
    __syncthreads();
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
    if (r < R.nranks) spin_until(R.my_flag + r, e, R.guard, kFaultAllreduce);
    __syncthreads();
    if (threadIdx.x == 0) NSB_STATS_ADD(R.stats, 0);
    if (threadIdx.x < count) {
        double s = 0.0;
        for (int q = 0; q < R.nranks; ++q) s += __ldcv(R.my_box + ((size_t)par * R.nranks + q) * kArSlots + threadIdx.x);
        vals[threadIdx.x] = s;
    }
    __syncthreads();
    if (threadIdx.x == 0) *R.seq = e;
}

// The tail of a kernel with a fused reduction, executed by all threads of its last block: thread 0 has already stored the rank's
// totals to vals[0..count).  `post` (the scalar step) runs in thread 0 once the totals are global.
template <class Post>
__device__ __forceinline__ void reduce_finish(const RedMode& M, double* vals, int count, Post post) {
    if (M.kind == kRedStore) return;
    if (M.kind == kRedPeer) p2p_allreduce_block(M.R, vals, count);
    if (threadIdx.x == 0) post();
}

// One block: all-reduce as a kernel of its own (set-up phases, Gram-Schmidt, the unfused paths).
template <class Post>
__global__ void k_p2p_allreduce(P2PReduceView R, double* vals, int count, Post post) {
    p2p_allreduce_block(R, vals, count);
    if (threadIdx.x == 0) post();
}

// ---- halo push from the producing kernel ---------------------------------------------------------------------------
// device view of one halo plan (p2p.cu)
struct HaloDev {
    int nnb;
    int64_t n_send, n_halo;
    const int* send_idx;     // owned entries to ship, neighbour by neighbour
    const int64_t* send_off; // nnb + 1 offsets into send_idx
    double* const* peer_land;   // per neighbour: where my entries land in its landing buffer
    u64* const* peer_ready;     // per neighbour: its ready flag for me
    u64* const* peer_consumed;  // per neighbour: its "neighbour consumed" word for me
    const double* my_land;
    int64_t land2;              // offset (doubles) of the second vector's area in every landing buffer
    const int64_t* peer_land2;  // per neighbour: the same offset in ITS landing buffer
    const u64* my_ready;        // written by the neighbours
    const u64* my_consumed;     // written by the neighbours
    u64* seq;
    unsigned int* ticket;
    u64* stats;   // NSB_SYNC_STATS builds; else null
    SpinGuard guard;
    int skip_publish;   // fault injection (tests): this rank never publishes "ready"
};

// The send lists of a plan as runs of consecutive owned rows: row i in [start, start + len) goes to neighbour nb, slot slot0 + i - start.
// tiles[t] != 0: some row of [256 t, 256 t + 256) is sent to a neighbour -- producers look at one byte per tile of rows and run the
// search through the runs only inside marked tiles (a test per row costs a memory-bound vector kernel ~5 %).
constexpr int kPushTileLog2 = 8;
struct PushOut {
    HaloDev H;
    int nruns;   // 0: nothing to push (single GPU, or the plan is not run-shaped: the push kernel is used)
    const unsigned char* tiles;
    int start[kPushRuns], len[kPushRuns], nb[kPushRuns], slot0[kPushRuns];
};

// Start of a pushing kernel, every block: the number of this exchange.  Block 0 tells the neighbours that what they pushed last
// time has been consumed here (true by stream order: the kernels that read the landing buffer came earlier on this stream).
__device__ __forceinline__ u64 push_begin(const PushOut& P) {
    const u64 e = *P.H.seq + 1;
    if (blockIdx.x == 0 && threadIdx.x < P.H.nnb) st_release_sys(P.H.peer_consumed[threadIdx.x], e - 1);
    return e;
}
// Before the first store of a block: the neighbours must have consumed what this rank pushed last time.  All threads call.
__device__ __forceinline__ void push_wait_consumed(const PushOut& P, u64 e) {
    if (threadIdx.x < P.H.nnb) spin_until(P.H.my_consumed + threadIdx.x, e - 1, P.H.guard, kFaultHaloConsumed);
    __syncthreads();
}
__device__ __forceinline__ bool push_touches(const PushOut& P, int64_t lo, int64_t hi) {   // rows [lo, hi) may meet a run
    unsigned int t = 0;
    for (int64_t k = lo >> kPushTileLog2; k <= (hi - 1) >> kPushTileLog2; ++k) t |= __ldg(P.tiles + k);
    return t != 0;
}
__device__ __forceinline__ bool push_tile(const PushOut& P, int64_t i) { return __ldg(P.tiles + (i >> kPushTileLog2)) != 0; }
// Row i of the produced vector(s): store into the landing buffers of the neighbours it is sent to (v2: the second vector of a
// dual exchange, area 2).
template <bool DUAL>
__device__ __forceinline__ void push_store(const PushOut& P, int64_t i, double v, double v2) {
    for (int k = 0; k < P.nruns; ++k) {
        const int64_t d = i - P.start[k];
        if (d >= 0 && d < P.len[k]) {
            double* land = P.H.peer_land[P.nb[k]];
            land[P.slot0[k] + d] = v;
            if (DUAL) land[P.H.peer_land2[P.nb[k]] + P.slot0[k] + d] = v2;
        }
    }
}
// End of a pushing kernel, all threads of every block: the last block to arrive publishes "ready" to every neighbour.
__device__ __forceinline__ void push_end(const PushOut& P, u64 e) {
    __threadfence_system();
    __syncthreads();
    __shared__ bool last_pusher;
    if (threadIdx.x == 0) last_pusher = atomicAdd(P.H.ticket, 1u) == gridDim.x - 1;
    __syncthreads();
    if (last_pusher) {
        __threadfence_system();
        if (threadIdx.x < P.H.nnb && !P.H.skip_publish) st_release_sys(P.H.peer_ready[threadIdx.x], e);
        if (threadIdx.x == 0) {
            *P.H.ticket = 0u;
            *P.H.seq = e;
        }
    }
}

P2PReduceView p2p_reduce_view(Context& c);   // p2p.cu
RedMode red_mode(Context& c);                // p2p.cu: kRedLocal / kRedPeer / kRedStore for this context
HaloIn p2p_halo_in(const nsb_halo_s* h, int area, const int* order, int n_interior);   // p2p.cu
PushOut p2p_push_out(const nsb_halo_s* h);   // p2p.cu: nruns == 0 when the producer cannot push for this plan
bool push_fused_ok(const Context& c, const nsb_mat_s* m);   // p2p.cu: consumer takes the halo in-kernel AND the plan is run-shaped

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
