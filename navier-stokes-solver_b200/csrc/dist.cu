// dist.cu -- row-partitioned mode: NCCL communicator (loaded at run time so single-GPU use has no
// NCCL dependency), interface-dof halo exchange before each SpMV and scalar all-reduce after each
// dot product.  The reference has no counterpart (single process).
#include <dlfcn.h>
#include <stdlib.h>

#include "common.cuh"

namespace nsb {

// minimal NCCL ABI (nccl.h, stable since 2.x)
typedef struct {
    char internal[128];
} NcclUniqueId;
typedef void* NcclComm;
enum { kNcclInt32 = 2, kNcclFloat64 = 8 };
enum { kNcclSum = 0 };

struct NcclApi {
    void* lib = nullptr;
    int (*GetUniqueId)(NcclUniqueId*) = nullptr;
    int (*CommInitRank)(NcclComm*, int, NcclUniqueId, int) = nullptr;
    int (*CommDestroy)(NcclComm) = nullptr;
    int (*AllReduce)(const void*, void*, size_t, int, int, NcclComm, cudaStream_t) = nullptr;
    int (*AllGather)(const void*, void*, size_t, int, NcclComm, cudaStream_t) = nullptr;
    int (*Send)(const void*, size_t, int, int, NcclComm, cudaStream_t) = nullptr;
    int (*Recv)(void*, size_t, int, int, NcclComm, cudaStream_t) = nullptr;
    int (*GroupStart)() = nullptr;
    int (*GroupEnd)() = nullptr;
    const char* (*GetErrorString)(int) = nullptr;
};

static NcclApi g_nccl;

static int load_nccl() {
    if (g_nccl.lib) return NSB_OK;
    const char* cands[4] = {getenv("NSB_NCCL_LIB"), "libnccl.so.2", "libnccl.so", nullptr};
    void* lib = nullptr;
    for (int i = 0; i < 3 && !lib; ++i)
        if (cands[i]) lib = dlopen(cands[i], RTLD_NOW | RTLD_GLOBAL);
    if (!lib) {
        set_error("NCCL library not loadable (libnccl.so.2): %s", dlerror());
        return NSB_ERR_NCCL;
    }
#define LOAD(field, name)                                                   \
    *(void**)(&g_nccl.field) = dlsym(lib, name);                            \
    if (!g_nccl.field) {                                                    \
        set_error("NCCL symbol %s missing", name);                          \
        return NSB_ERR_NCCL;                                                \
    }
    LOAD(GetUniqueId, "ncclGetUniqueId");
    LOAD(CommInitRank, "ncclCommInitRank");
    LOAD(CommDestroy, "ncclCommDestroy");
    LOAD(AllReduce, "ncclAllReduce");
    LOAD(AllGather, "ncclAllGather");
    LOAD(Send, "ncclSend");
    LOAD(Recv, "ncclRecv");
    LOAD(GroupStart, "ncclGroupStart");
    LOAD(GroupEnd, "ncclGroupEnd");
    LOAD(GetErrorString, "ncclGetErrorString");
#undef LOAD
    g_nccl.lib = lib;
    return NSB_OK;
}

#define NSB_NCCL(call)                                                                              \
    do {                                                                                            \
        int r__ = (call);                                                                           \
        if (r__ != 0) {                                                                             \
            set_error("%s:%d: %s -> %s", __FILE__, __LINE__, #call, g_nccl.GetErrorString(r__));    \
            return NSB_ERR_NCCL;                                                                    \
        }                                                                                           \
    } while (0)

int dist_init(Context& c, int rank, int nranks, const char* uid) {
    NSB_CHECK(load_nccl());
    NcclUniqueId id;
    memcpy(id.internal, uid, 128);
    NcclComm comm = nullptr;
    NSB_NCCL(g_nccl.CommInitRank(&comm, nranks, id, rank));
    c.nccl_comm = comm;
    c.nccl = &g_nccl;
    c.rank = rank;
    c.nranks = nranks;
    const char* env = getenv("NSB_P2P");
    if (!(env && atoi(env) == 0)) NSB_CHECK(p2p_reduce_init(c));   // peer-memory scalar all-reduce when IPC works
    return NSB_OK;
}

int nccl_allgather_bytes(Context& c, const void* send_dev, void* recv_dev, size_t bytes_per_rank) {
    NSB_NCCL(c.nccl->AllGather(send_dev, recv_dev, bytes_per_rank, 0 /* ncclInt8 */, (NcclComm)c.nccl_comm, c.stream));
    NSB_CUDA(cudaStreamSynchronize(c.stream));
    return NSB_OK;
}

void dist_destroy(Context& c) {
    p2p_reduce_destroy(c);
    if (c.nccl_comm && c.nccl) c.nccl->CommDestroy((NcclComm)c.nccl_comm);
    c.nccl_comm = nullptr;
}

int allreduce_sum(Context& c, double* dev, int count) {
    if (c.nranks <= 1) return NSB_OK;
    if (c.p2p && count <= 8) return p2p_allreduce(c, dev, count);
    NSB_NCCL(c.nccl->AllReduce(dev, dev, (size_t)count, kNcclFloat64, kNcclSum, (NcclComm)c.nccl_comm, c.stream));
    return NSB_OK;
}

__global__ void __launch_bounds__(kBlock) k_pack(double* __restrict__ buf, const double* __restrict__ x,
                                                 const int* __restrict__ idx, int64_t n) {
    for (int64_t i = (int64_t)blockIdx.x * blockDim.x + threadIdx.x; i < n; i += (int64_t)gridDim.x * blockDim.x)
        buf[i] = x[idx[i]];
}

int halo_exchange(Context& c, nsb_halo_s* h, double* x) {
    if (!h || c.nranks <= 1) return NSB_OK;
    if (h->p2p) return p2p_halo_exchange(c, h, x);
    if (h->n_send > 0) {
        k_pack<<<vec_grid(h->n_send, c.num_sms), kBlock, 0, c.stream>>>(h->send_buf.p, x, h->send_idx.p, h->n_send);
        c.launches++;
        NSB_CUDA(cudaGetLastError());
    }
    NSB_NCCL(c.nccl->GroupStart());
    int64_t so = 0, ro = 0;
    for (size_t k = 0; k < h->neighbors.size(); ++k) {
        if (h->send_counts[k] > 0)
            NSB_NCCL(c.nccl->Send(h->send_buf.p + so, (size_t)h->send_counts[k], kNcclFloat64, h->neighbors[k],
                                  (NcclComm)c.nccl_comm, c.stream));
        if (h->recv_counts[k] > 0)
            NSB_NCCL(c.nccl->Recv(x + h->n_owned + ro, (size_t)h->recv_counts[k], kNcclFloat64, h->neighbors[k],
                                  (NcclComm)c.nccl_comm, c.stream));
        so += h->send_counts[k];
        ro += h->recv_counts[k];
    }
    NSB_NCCL(c.nccl->GroupEnd());
    return NSB_OK;
}

int halo_exchange2(Context& c, nsb_halo_s* h, double* x1, double* x2) {
    if (!h || c.nranks <= 1) return NSB_OK;
    if (h->p2p) return p2p_halo_exchange(c, h, x1, x2);
    NSB_CHECK(halo_exchange(c, h, x1));
    return halo_exchange(c, h, x2);
}

}  // namespace nsb

nsb_halo_s::~nsb_halo_s() { nsb::p2p_halo_destroy(this); }

using namespace nsb;

extern "C" {

int nsb_nccl_unique_id(char out[128]) {
    NSB_REQUIRE(out, "nsb_nccl_unique_id: out is NULL");
    NSB_CHECK(load_nccl());
    NcclUniqueId id;
    NSB_NCCL(g_nccl.GetUniqueId(&id));
    memcpy(out, id.internal, 128);
    return NSB_OK;
}

int nsb_halo_create(nsb_ctx ctx, int64_t n_owned, int64_t n_halo, int n_neighbors, const int32_t* neighbors,
                    const int64_t* send_counts, const int32_t* send_idx, const int64_t* recv_counts, nsb_halo* out) {
    NSB_REQUIRE(ctx && out, "nsb_halo_create: NULL argument");
    NSB_REQUIRE(n_owned >= 0 && n_halo >= 0 && n_neighbors >= 0, "nsb_halo_create: negative size");
    nsb_halo_s* h = new nsb_halo_s();
    h->ctx = &ctx->c;
    h->n_owned = n_owned;
    h->n_halo = n_halo;
    int64_t ns = 0, nr = 0;
    for (int k = 0; k < n_neighbors; ++k) {
        h->neighbors.push_back(neighbors[k]);
        h->send_counts.push_back(send_counts[k]);
        h->recv_counts.push_back(recv_counts[k]);
        ns += send_counts[k];
        nr += recv_counts[k];
    }
    if (nr != n_halo) {
        delete h;
        set_error("nsb_halo_create: recv_counts sum %lld != n_halo %lld", (long long)nr, (long long)n_halo);
        return NSB_ERR_INVALID;
    }
    h->n_send = ns;
    int r = h->send_idx.alloc(ns > 0 ? ns : 1);
    if (r == NSB_OK) r = h->send_buf.alloc(ns > 0 ? ns : 1);
    if (r == NSB_OK && ns > 0) {
        cudaError_t e = cudaMemcpy(h->send_idx.p, send_idx, sizeof(int32_t) * (size_t)ns, cudaMemcpyDefault);
        if (e != cudaSuccess) {
            set_error("nsb_halo_create: copy failed: %s", cudaGetErrorString(e));
            r = NSB_ERR_CUDA;
        }
    }
    if (r == NSB_OK && ctx->c.nranks > 1 && ctx->c.p2p) r = p2p_halo_init(ctx->c, h);   // collective
    if (r != NSB_OK) {
        delete h;
        return r;
    }
    *out = h;
    return NSB_OK;
}

int nsb_halo_destroy(nsb_halo h) {
    delete h;
    return NSB_OK;
}

int nsb_mat_set_halo(nsb_mat m, nsb_halo h) {
    NSB_REQUIRE(m, "nsb_mat_set_halo: NULL matrix");
    if (h) NSB_REQUIRE(h->n_owned + h->n_halo == m->cols, "nsb_mat_set_halo: n_owned+n_halo = %lld but matrix has %lld columns",
                       (long long)(h->n_owned + h->n_halo), (long long)m->cols);
    m->halo = h;
    return m->ctx ? build_halo_order(*m->ctx, m) : NSB_OK;
}

}  // extern "C"
