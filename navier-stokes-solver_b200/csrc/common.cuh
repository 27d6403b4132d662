// common.cuh -- internal declarations shared by the translation units of libnsb200.so
// (context, device buffers, error handling, deterministic block/grid reductions).
#pragma once

#include <cuda_runtime.h>
#include <stdint.h>
#include <stdio.h>
#include <string.h>

#include <memory>
#include <string>
#include <utility>
#include <vector>

#include <nvtx3/nvToolsExt.h>

#include "../../include/nsb200.h"

namespace nsb {

// ---------------------------------------------------------------------------------------------
// errors
// ---------------------------------------------------------------------------------------------
void set_error(const char* fmt, ...);
extern thread_local char g_err[1024];

#define NSB_CUDA(call)                                                                          \
    do {                                                                                        \
        cudaError_t e__ = (call);                                                               \
        if (e__ != cudaSuccess) {                                                               \
            nsb::set_error("%s:%d: %s -> %s", __FILE__, __LINE__, #call, cudaGetErrorString(e__)); \
            return NSB_ERR_CUDA;                                                                \
        }                                                                                       \
    } while (0)

#define NSB_CHECK(call)                  \
    do {                                 \
        int r__ = (call);                \
        if (r__ != NSB_OK) return r__;   \
    } while (0)

#define NSB_REQUIRE(cond, ...)           \
    do {                                 \
        if (!(cond)) {                   \
            nsb::set_error(__VA_ARGS__); \
            return NSB_ERR_INVALID;      \
        }                                \
    } while (0)

// ---------------------------------------------------------------------------------------------
// tracing: one NVTX range per solver phase (the reference brackets the same phases with NGSolve Timers: bramblepasciak_new.py
// :111-116,:194-196, bramble_pasciak_cg.py:111-113, orthonormalization.py:6-14, heat.py:85-128).  Header-only NVTX v3: a no-op unless
// a profiler (nsys, ncu --nvtx) is attached.
// ---------------------------------------------------------------------------------------------
struct NvtxRange {
    explicit NvtxRange(const char* name) { nvtxRangePushA(name); }
    NvtxRange(const NvtxRange&) = delete;
    NvtxRange& operator=(const NvtxRange&) = delete;
    ~NvtxRange() { nvtxRangePop(); }
};

// ---------------------------------------------------------------------------------------------
// launch geometry.  Vector kernels are persistent grid-stride loops: a fixed grid (a multiple of
// the SM count) keeps every reduction tree identical from run to run for a given length.
// ---------------------------------------------------------------------------------------------
constexpr int kBlock = 256;          // threads per block of the vector kernels
constexpr int kMaxBlocks = 148 * 8;  // 148 SMs x 8 resident 256-thread blocks
constexpr int kMaxRed = 8;           // doubles reduced per kernel at most
// SELL-LX SpMV (spmv.cuh, sell.cu): shared-memory doubles a chunk may stage (distinct columns + R row
// sums; 36 KB per block), target nonzeros per chunk
constexpr int kSellStage = 4600;
constexpr int kSellStageMax = 7300;   // relaxed budget for chunks of long rows (sell.cu)
constexpr double kSellTargetNnz = 24000.0;

inline int vec_grid(int64_t n, int num_sms) {
    int64_t want = (n + (int64_t)kBlock * 2 - 1) / ((int64_t)kBlock * 2);
    int64_t cap = (int64_t)num_sms * 8;
    if (want < 1) want = 1;
    return (int)(want < cap ? want : cap);
}

// ---------------------------------------------------------------------------------------------
// context
// ---------------------------------------------------------------------------------------------
struct NcclApi;    // dist.cu
struct P2PReduce;  // p2p.cu
struct P2PHalo;    // p2p.cu

struct Context {
    int device = 0;
    int num_sms = 148;
    cudaStream_t stream = nullptr;
    int64_t launches = 0;
    bool use_graphs = true;
    // programmatic dependent launch (NSB_PDL, default OFF -- measured 0..+0.8 % / -12 %): the kernels of an iteration loop are launched with
    // cudaLaunchAttributeProgrammaticStreamSerialization, so the next kernel's blocks become resident while the
    // previous kernel drains, fetch their immutable matrix structure and park in griddepcontrol.wait
    bool pdl = false;
    int pdl_mask = 3;    // NSB_PDL = 1 | 2 | 3: bit 0 SpMV kernels, bit 1 vector kernels launched with the attribute
    // SELL kernel options (NSB_SELL_FLAGS): 1 trigger dependents at entry, 2 prefetch the first chunk's structure
    // before the dependency wait (both only matter under NSB_PDL), 4 epilogue with the operand loads of two rows per
    // thread in flight, 8 x staged with cp.async, 16 the next chunk staged under the current chunk's epilogue
    int sell_flags = 31;
    // reduction scratch: per-block partials and the arrival ticket of the last-block pattern
    double* partials = nullptr;  // kMaxRed * max grid
    double* chain = nullptr;     // kMaxRed: totals handed from one row panel's launch to the next (spmv.cuh: ChainOp)
    unsigned int* ticket = nullptr;
    int max_grid = 0;
    // pinned host mailbox for scalar read-back
    double* h_mail = nullptr;
    double* d_mail = nullptr;
    // distributed
    int rank = 0, nranks = 1;
    void* nccl_comm = nullptr;
    NcclApi* nccl = nullptr;
    P2PReduce* p2p = nullptr;   // peer-memory scalar all-reduce (p2p.cu); null: NCCL all-reduce
    // -DNSB_SYNC_STATS builds only (tools/sync_stats.sh): nanoseconds spent polling peer flags and poll counts,
    // [0,1] all-reduce  [2,3] push waiting for "consumed"  [4,5] SpMV block 0 waiting for the halo  [6,7] wait/copy kernel
    unsigned long long* sync_stats = nullptr;
    // failure detection of the peer-memory protocol: a wait on a peer flag that exceeds spin_timeout_ns raises *fault (device
    // word) and is abandoned; the solver call then returns NSB_ERR_NCCL (check_fault).  NSB_SPIN_TIMEOUT_MS, default 30 s, 0 = never
    int* fault = nullptr;
    unsigned long long spin_timeout_ns = 30000000000ull;
    int fault_skip_publish = 0;   // NSB_FAULT_SKIP_PUBLISH=<rank>: that rank never publishes "halo ready" (tests of the bound)
    // launch geometry of the SELL kernels per (instantiation, shared-memory size): blocks per SM (spmv.cuh)
    std::vector<std::pair<std::pair<const void*, size_t>, int>> sell_occupancy;
};

// device buffer with RAII (host-side owner).  alloc(): cudaMalloc, for long-lived storage (operators, IPC
// blocks).  allocs(): stream-ordered allocation from the context's memory pool, for the work vectors of a
// solve -- a solve allocates ~20 vectors and cudaMalloc/cudaFree of 150 MB blocks cost ~30 ms per call.
template <typename T>
struct DevBuf {
    T* p = nullptr;
    int64_t n = 0;
    cudaStream_t pool_stream = nullptr;
    bool pooled = false;
    DevBuf() = default;
    DevBuf(const DevBuf&) = delete;
    DevBuf& operator=(const DevBuf&) = delete;
    ~DevBuf() { release(); }
    void release() {
        if (p) {
            if (pooled) cudaFreeAsync(p, pool_stream);
            else cudaFree(p);
        }
        p = nullptr;
        n = 0;
        pooled = false;
    }
    int alloc(int64_t count) {
        release();
        n = count;
        if (count <= 0) return NSB_OK;
        cudaError_t e = cudaMalloc((void**)&p, sizeof(T) * (size_t)count);
        if (e != cudaSuccess) {
            p = nullptr;
            cudaGetLastError();   // reported here: do not leave it behind for the next library call (cub peeks at the last error)
            set_error("cudaMalloc(%lld bytes) failed: %s", (long long)(sizeof(T) * (size_t)count), cudaGetErrorString(e));
            return NSB_ERR_CUDA;
        }
        return NSB_OK;
    }
    int allocs(Context& c, int64_t count) {
        release();
        n = count;
        if (count <= 0) return NSB_OK;
        cudaError_t e = cudaMallocAsync((void**)&p, sizeof(T) * (size_t)count, c.stream);
        if (e != cudaSuccess) {
            p = nullptr;
            cudaGetLastError();
            set_error("cudaMallocAsync(%lld bytes) failed: %s", (long long)(sizeof(T) * (size_t)count), cudaGetErrorString(e));
            return NSB_ERR_CUDA;
        }
        pooled = true;
        pool_stream = c.stream;
        return NSB_OK;
    }
};

// SELL-LX storage of one row panel [row0, row0 + nrows) of an operator (sell.cu, spmv.cuh)
struct SellData {
    int64_t row0 = 0, nrows = 0;
    int sell_R = 0, sell_log2R = 0, sell_T = 1, sell_nchunks = 0, sell_max_uniq = 0, sell_cap = 0;
    int64_t sell_total_uniq = 0, sell_stored = 0;   // distinct-column list length; stored (padded) nonzeros
    nsb::DevBuf<double> sell_vals;            // pair-blocks of 32 x 2 values
    nsb::DevBuf<unsigned short> sell_lcol;    // same layout: index into the chunk's distinct-column list
    nsb::DevBuf<unsigned int> sell_off;       // per slice (+1), in pair-blocks
    nsb::DevBuf<unsigned short> sell_perm;    // nchunks * R: sorted position -> row inside the chunk
    nsb::DevBuf<int32_t> sell_uptr;           // nchunks + 1
    nsb::DevBuf<int32_t> sell_ucol;           // distinct columns, chunk by chunk, ascending
    // the same lists as runs of consecutive columns (used when the mean run is >= 8): first column, position
    int sell_nruns = 0;
    nsb::DevBuf<int32_t> sell_rstart, sell_rpos, sell_rptr;   // nruns, nruns + 1, nchunks + 1
    // row-partitioned operators (sell.cu: build_halo_order): chunk visiting order with the chunks that touch no halo
    // column first, so a SpMV can start before the neighbours' interface entries have landed (p2p.cuh: HaloIn)
    nsb::DevBuf<int32_t> halo_order;
    int halo_n_interior = -1;   // -1: no order built
};

}  // namespace nsb

// ---------------------------------------------------------------------------------------------
// opaque handle bodies
// ---------------------------------------------------------------------------------------------
struct nsb_ctx_s {
    nsb::Context c;
};

struct nsb_halo_s {
    nsb::Context* ctx = nullptr;
    int64_t n_owned = 0, n_halo = 0;
    std::vector<int> neighbors;
    std::vector<int64_t> send_counts, recv_counts;
    int64_t n_send = 0;
    nsb::DevBuf<int32_t> send_idx;  // owned entries to ship, neighbour by neighbour
    nsb::DevBuf<double> send_buf;
    nsb::P2PHalo* p2p = nullptr;    // peer-memory push exchange (p2p.cu); null: pack + NCCL send/recv
    ~nsb_halo_s();
};

struct nsb_mat_s {
    nsb::Context* ctx = nullptr;
    int64_t rows = 0, cols = 0, nnz = 0;
    nsb::DevBuf<int32_t> rowptr;  // rows + 1
    nsb::DevBuf<int32_t> colidx;  // nnz
    nsb::DevBuf<double> values;   // nnz
    int threads_per_row = 8;      // sub-warp width of the CSR-vector kernel, chosen from nnz/rows
    // SELL-LX companion (sell.cu): one SellData per row panel (a saddle-point operator has two: velocity rows
    // [A B^T] and pressure rows [B 0], whose row lengths differ by a factor of five); empty when absent (small or
    // too scattered operators: the CSR-vector kernel runs)
    std::vector<std::unique_ptr<nsb::SellData>> panels;
    std::vector<int64_t> panel_splits;   // row boundaries requested before build_sell (default: one panel)
    nsb_halo_s* halo = nullptr;   // not owned
    // saddle composition bookkeeping (rows of the velocity block; 0 for a plain matrix)
    int64_t n_u = 0, n_p = 0;
};

struct nsb_pre_s {
    nsb::Context* ctx = nullptr;
    int64_t n = 0;
    int kind = 0;           // 0: diagonal (y = d .* x)   1: block smoother   2: auxiliary-space composite   3: block-diagonal of two (blockpre.cu)
    nsb::DevBuf<double> d;  // diagonal preconditioner: y = d .* x
    // block smoother, a.mat.CreateBlockSmoother(blocks) (templates/NavierStokesSIMPLE_iterative.py:360-373): disjoint dof
    // blocks listed colour by colour (blocks of one colour are not coupled by the matrix), dense inverse per block
    nsb_mat_s* mat = nullptr;              // not owned
    int64_t nblocks = 0;
    int ncolors = 0, max_block = 0, gs = 0; // gs: nsb_pre_apply runs the symmetric block Gauss-Seidel cycle instead of block Jacobi
    std::vector<int64_t> color_ptr;        // host copy: blocks of colour c are [color_ptr[c], color_ptr[c+1])
    nsb::DevBuf<int64_t> block_ptr;        // nblocks + 1, into block_dofs
    nsb::DevBuf<int32_t> block_dofs;
    nsb::DevBuf<int64_t> inv_ptr;          // nblocks + 1, into inv (block b: size^2 entries, row-major)
    nsb::DevBuf<double> inv;
    nsb::DevBuf<double> ybuf;              // row-partitioned smoother only: y with a halo tail (mat->cols entries) for the GS sweeps
    // kind 2: auxiliary-space preconditioner MypreA (:364-383) = block smoother + transform coarse transform^T.
    // Nothing below is owned.  gs selects the multiplicative (:375-381) or the additive (:383) variant.
    nsb_pre_s* smoother = nullptr;         // kind 1
    nsb_pre_s* coarse = nullptr;           // any kind, size = transform columns
    nsb_mat_s *transform = nullptr, *transform_t = nullptr;
    nsb::DevBuf<double> aux_t, aux_c, aux_z;   // scratch: n, n_coarse, n_coarse (aux_t / aux_z with a halo tail when row-partitioned)
    // kind 3: BlockMatrix([[blk0, None], [None, blk1]]) of two preconditioners of any kind (run.py:46), not owned
    nsb_pre_s *blk0 = nullptr, *blk1 = nullptr;
};

namespace nsb {

// ---------------------------------------------------------------------------------------------
// device helpers
// ---------------------------------------------------------------------------------------------
#ifdef __CUDACC__

// Programmatic dependent launch (sm_90+).  pdl_trigger(): the next kernel in the stream may be scheduled as soon as
// every block of this grid has executed it (its blocks then fill the SMs this grid frees while it drains).
// pdl_wait(): returns when the preceding grid has completed and its writes are visible; everything a kernel reads or
// writes that an earlier kernel may have written comes after it.  EVERY block of a kernel launched through
// launch_k() executes pdl_wait() before it exits, so "this grid completed" implies "its predecessor completed".
// Both are no-ops in a kernel launched without the attribute.
__device__ __forceinline__ void pdl_trigger() { asm volatile("griddepcontrol.launch_dependents;" ::: "memory"); }
// The fence behind the wait compiles to MEMBAR + CCTL.IVALL: a block that became resident early shares its SM's L1 with
// blocks of the two previous grids, which may have pulled in lines (plain or ld.global.nc loads) that the previous
// grid overwrote afterwards; the invalidate makes the loads behind the wait miss L1 like they do after an ordinary
// kernel boundary.
__device__ __forceinline__ void pdl_wait() { asm volatile("griddepcontrol.wait;\n\tfence.acq_rel.gpu;" ::: "memory"); }
__device__ __forceinline__ void pdl_enter() {
    pdl_trigger();
    pdl_wait();
}
__device__ __forceinline__ void prefetch_l2(const void* p) { asm volatile("prefetch.global.L2 [%0];" ::"l"(p)); }

__device__ __forceinline__ double warp_sum(double v) {
#pragma unroll
    for (int o = 16; o > 0; o >>= 1) v += __shfl_xor_sync(0xffffffffu, v, o);
    return v;
}

// Block-wide sum of NS values with a fixed tree (warp shuffle, then one warp over the warp sums).
// Result valid in thread 0.
template <int NS>
__device__ __forceinline__ void block_sum(double (&v)[NS]) {
    __shared__ double sm[NS][32];
    const int lane = threadIdx.x & 31, wid = threadIdx.x >> 5;
    const int nw = (blockDim.x + 31) >> 5;
#pragma unroll
    for (int k = 0; k < NS; ++k) {
        double s = warp_sum(v[k]);
        if (lane == 0) sm[k][wid] = s;
    }
    __syncthreads();
    if (wid == 0) {
#pragma unroll
        for (int k = 0; k < NS; ++k) {
            double s = lane < nw ? sm[k][lane] : 0.0;
            v[k] = warp_sum(s);
        }
    }
    __syncthreads();
}

// Grid-wide deterministic reduction, "last block finishes" pattern.  Every block publishes its NS
// block sums, takes a ticket, and the block that draws the last ticket sums the partials of all
// blocks in a fixed order.  Returns true in ALL threads of that last block, with the totals in
// v[] valid in thread 0.  The ticket is reset for the next kernel.
template <int NS>
__device__ __forceinline__ bool grid_sum(double (&v)[NS], double* partials, unsigned int* ticket) {
    block_sum<NS>(v);
    __shared__ bool is_last;
    if (threadIdx.x == 0) {
#pragma unroll
        for (int k = 0; k < NS; ++k) partials[(size_t)k * gridDim.x + blockIdx.x] = v[k];
        __threadfence();
        unsigned int t = atomicAdd(ticket, 1u);
        is_last = (t == gridDim.x - 1);
    }
    __syncthreads();
    if (!is_last) return false;
    __threadfence();
#pragma unroll
    for (int k = 0; k < NS; ++k) {
        double s = 0.0;
        for (unsigned int b = threadIdx.x; b < gridDim.x; b += blockDim.x)
            s += __ldcg(&partials[(size_t)k * gridDim.x + b]);
        v[k] = s;
    }
    block_sum<NS>(v);
    if (threadIdx.x == 0) *ticket = 0u;
    return true;
}

#endif  // __CUDACC__

#ifdef __CUDACC__
// Launch of a kernel that takes part in the programmatic-dependent-launch protocol above (it calls pdl_wait()
// before touching mutable data).  With c.pdl off this is an ordinary stream-ordered launch.
template <class... KArgs, class... Args>
inline cudaError_t launch_k(Context& c, int kind, void (*kern)(KArgs...), int grid, int block, size_t smem, Args&&... args) {
    cudaLaunchConfig_t cfg = {};
    cfg.gridDim = dim3((unsigned)grid, 1, 1);
    cfg.blockDim = dim3((unsigned)block, 1, 1);
    cfg.dynamicSmemBytes = smem;
    cfg.stream = c.stream;
    cudaLaunchAttribute at[1];
    at[0].id = cudaLaunchAttributeProgrammaticStreamSerialization;
    at[0].val.programmaticStreamSerializationAllowed = 1;
    cfg.attrs = at;
    cfg.numAttrs = (c.pdl && (c.pdl_mask & kind)) ? 1 : 0;
    return cudaLaunchKernelEx(&cfg, kern, static_cast<KArgs>(args)...);
}
#endif

// ---------------------------------------------------------------------------------------------
// host-side utilities implemented in context.cu / vecops.cu
// ---------------------------------------------------------------------------------------------
int copy_in(Context& c, double* dst_dev, const double* src_any, int64_t n);   // host|device -> device, async on stream
int copy_out(Context& c, double* dst_any, const double* src_dev, int64_t n);  // device -> host|device, async on stream
bool is_device_ptr(const void* p);

// generic (unfused) vector ops used by set-up phases
int vec_set(Context& c, double* y, double a, int64_t n);
int vec_lincomb(Context& c, double* y, double a, const double* x1, double b, const double* x2, double cc,
                const double* x3, int64_t n);                               // y = a x1 + b x2 + cc x3 (x2,x3 may be null)
int vec_diagmul(Context& c, double* y, double s, const double* d, const double* x, int64_t n);  // y = s d.*x (d null: y = s x)
int vec_dot(Context& c, const double* x, const double* y, int64_t n, double* out_dev);          // *out_dev = <x,y> (+ allreduce when distributed)
int allreduce_sum(Context& c, double* dev, int count);                       // no-op for nranks == 1
int halo_exchange(Context& c, nsb_halo_s* h, double* x);                     // fills x[n_owned .. n_owned+n_halo)
// peer-memory (NVLink) paths, p2p.cu
int p2p_reduce_init(Context& c);                                              // collective; leaves c.p2p null when IPC is unavailable
void p2p_reduce_destroy(Context& c);
int p2p_allreduce(Context& c, double* dev, int count);
int p2p_halo_init(Context& c, nsb_halo_s* h);                                 // collective; leaves h->p2p null on failure
void p2p_halo_destroy(nsb_halo_s* h);
int p2p_halo_exchange(Context& c, nsb_halo_s* h, double* x, double* x2 = nullptr);   // x2: a second vector in the same exchange
int halo_exchange2(Context& c, nsb_halo_s* h, double* x1, double* x2);             // dist.cu
int nccl_allgather_bytes(Context& c, const void* send_dev, void* recv_dev, size_t bytes_per_rank);   // dist.cu

// Host side of an iteration loop whose stop test runs on the device.  `batch()` enqueues `batch_iters` iterations
// (one graph launch, or plain launches); batches are enqueued until the device-side stop flag is seen set or
// `maxsteps` iterations are in the stream.  The flag of batch k is read only after batch k+1 has been enqueued:
// the kernels of an iteration are no-ops once the flag is set, so running one batch ahead costs a few empty
// launches at the end of a solve and the device never idles on the host round trip in between.  In row-partitioned
// runs every rank reads the same flag values in the same order, so all ranks enqueue the same number of batches.
template <class Batch>
int drive_batches(Context& c, const int* stop_dev, int maxsteps, int batch_iters, Batch batch) {
    volatile int* h_flags = (volatile int*)c.h_mail;   // pinned
    cudaEvent_t ev[2] = {nullptr, nullptr};
    NSB_CUDA(cudaEventCreateWithFlags(&ev[0], cudaEventDisableTiming));
    NSB_CUDA(cudaEventCreateWithFlags(&ev[1], cudaEventDisableTiming));
    int rc = NSB_OK, launched = 0, slot = 0;
    bool stop = false, have_prev = false;
    cudaError_t e = cudaSuccess;
    while (rc == NSB_OK && e == cudaSuccess && !stop && launched < maxsteps) {
        rc = batch();
        if (rc != NSB_OK) break;
        launched += batch_iters;
        h_flags[slot] = 0;
        h_flags[2 + slot] = 0;
        e = cudaMemcpyAsync((void*)&h_flags[slot], stop_dev, sizeof(int), cudaMemcpyDeviceToHost, c.stream);
        // row-partitioned: a peer exchange that timed out (p2p.cuh: spin_until) ends the loop as well; the caller reports it
        if (e == cudaSuccess && c.nranks > 1 && c.fault)
            e = cudaMemcpyAsync((void*)&h_flags[2 + slot], c.fault, sizeof(int), cudaMemcpyDeviceToHost, c.stream);
        if (e == cudaSuccess) e = cudaEventRecord(ev[slot], c.stream);
        if (e == cudaSuccess && have_prev) {
            e = cudaEventSynchronize(ev[slot ^ 1]);
            stop = h_flags[slot ^ 1] != 0 || h_flags[2 + (slot ^ 1)] != 0;
        }
        have_prev = true;
        slot ^= 1;
    }
    if (e == cudaSuccess) e = cudaStreamSynchronize(c.stream);
    cudaEventDestroy(ev[0]);
    cudaEventDestroy(ev[1]);
    if (rc == NSB_OK && e != cudaSuccess) {
        set_error("iteration loop: %s", cudaGetErrorString(e));
        rc = NSB_ERR_CUDA;
    }
    return rc;
}

// two timing events that are destroyed on every exit path of a solver
struct EventPair {
    cudaEvent_t e0 = nullptr, e1 = nullptr;
    EventPair() = default;
    EventPair(const EventPair&) = delete;
    EventPair& operator=(const EventPair&) = delete;
    int create() {
        NSB_CUDA(cudaEventCreate(&e0));
        NSB_CUDA(cudaEventCreate(&e1));
        return NSB_OK;
    }
    ~EventPair() {
        if (e0) cudaEventDestroy(e0);
        if (e1) cudaEventDestroy(e1);
    }
};

// Capture `n` calls of `one()` on the context's stream into an executable graph.  On any failure nothing is leaked and *exec
// stays null.  *launches = kernels the captured batch launches (c.launches is left as it was).
template <class One>
int capture_batch(Context& c, int n, One one, cudaGraphExec_t* exec, int64_t* launches, const char* who) {
    *exec = nullptr;
    const int64_t before = c.launches;
    cudaGraph_t graph = nullptr;
    NSB_CUDA(cudaStreamBeginCapture(c.stream, cudaStreamCaptureModeThreadLocal));
    int rc = NSB_OK;
    for (int j = 0; j < n && rc == NSB_OK; ++j) rc = one(j);
    cudaError_t ce = cudaStreamEndCapture(c.stream, &graph);
    *launches = c.launches - before;
    c.launches = before;
    if (rc == NSB_OK && ce != cudaSuccess) {
        set_error("%s: graph capture failed: %s", who, cudaGetErrorString(ce));
        rc = NSB_ERR_CUDA;
    }
    if (rc == NSB_OK && (ce = cudaGraphInstantiate(exec, graph, 0)) != cudaSuccess) {
        set_error("%s: graph instantiation failed: %s", who, cudaGetErrorString(ce));
        *exec = nullptr;
        rc = NSB_ERR_CUDA;
    }
    if (graph) cudaGraphDestroy(graph);
    return rc;
}

struct GraphExecHolder {
    cudaGraphExec_t exec = nullptr;
    GraphExecHolder() = default;
    GraphExecHolder(const GraphExecHolder&) = delete;
    GraphExecHolder& operator=(const GraphExecHolder&) = delete;
    ~GraphExecHolder() {
        if (exec) cudaGraphExecDestroy(exec);
    }
};

int build_sell(Context& c, nsb_mat_s* m);   // sell.cu
int build_halo_order(Context& c, nsb_mat_s* m);   // sell.cu: (re)build the interior-first chunk order of every panel for m->halo
// context.cu: NSB_OK, or NSB_ERR_NCCL (fault word cleared, error text set) when a peer exchange timed out since the last check
int check_fault(Context& c, const char* who);
// exchange steps of the fused path (p2p.cu): push only; the consuming SpMV waits inside the kernel
int p2p_halo_push(Context& c, nsb_halo_s* h, double* x, double* x2 = nullptr);
bool halo_fused_ok(const Context& c, const nsb_mat_s* m);   // p2p.cu: operator and exchange qualify for the in-kernel wait
// y = P x for any preconditioner kind, device vectors (blockpre.cu); y must not alias x
int pre_apply_device(Context& c, nsb_pre_s* p, const double* x, double* y);

// plain SpMV launch (spmv.cu): y = M x on the context stream; x must have room for the halo tail
int spmv_plain(Context& c, const nsb_mat_s* m, const double* x, double* y);

}  // namespace nsb
