// context.cu -- context lifetime, error text, host<->device staging.
#include <stdarg.h>
#include <stdlib.h>

#include "common.cuh"

namespace nsb {

thread_local char g_err[1024] = {0};

void set_error(const char* fmt, ...) {
    va_list ap;
    va_start(ap, fmt);
    vsnprintf(g_err, sizeof(g_err), fmt, ap);
    va_end(ap);
}

bool is_device_ptr(const void* p) {
    cudaPointerAttributes a;
    cudaError_t e = cudaPointerGetAttributes(&a, p);
    if (e != cudaSuccess) {
        cudaGetLastError();
        return false;
    }
    return a.type == cudaMemoryTypeDevice || a.type == cudaMemoryTypeManaged;
}

int copy_in(Context& c, double* dst, const double* src, int64_t n) {
    if (n <= 0) return NSB_OK;
    NSB_CUDA(cudaMemcpyAsync(dst, src, sizeof(double) * (size_t)n, cudaMemcpyDefault, c.stream));
    return NSB_OK;
}

int copy_out(Context& c, double* dst, const double* src, int64_t n) {
    if (n <= 0) return NSB_OK;
    NSB_CUDA(cudaMemcpyAsync(dst, src, sizeof(double) * (size_t)n, cudaMemcpyDefault, c.stream));
    return NSB_OK;
}

int ctx_init(Context& c, int device) {
    c.device = device;
    NSB_CUDA(cudaSetDevice(device));
    cudaDeviceProp prop;
    NSB_CUDA(cudaGetDeviceProperties(&prop, device));
    c.num_sms = prop.multiProcessorCount;
    {
        const char* e = getenv("NSB_PDL");
        c.pdl = (e ? atoi(e) != 0 : false) && prop.major >= 9;
        if (e && atoi(e) > 0) c.pdl_mask = atoi(e) & 3;
        const char* f = getenv("NSB_SELL_FLAGS");
        if (f) c.sell_flags = atoi(f);
        const char* g = getenv("NSB_GRAPHS");
        if (g) c.use_graphs = atoi(g) != 0;
        const char* t = getenv("NSB_SPIN_TIMEOUT_MS");
        if (t) c.spin_timeout_ns = (unsigned long long)(atof(t) * 1e6);
    }
    NSB_CUDA(cudaStreamCreateWithFlags(&c.stream, cudaStreamNonBlocking));
    {   // keep freed work vectors in the device's default pool between solves
        cudaMemPool_t pool;
        NSB_CUDA(cudaDeviceGetDefaultMemPool(&pool, device));
        unsigned long long keep = ~0ull;
        NSB_CUDA(cudaMemPoolSetAttribute(pool, cudaMemPoolAttrReleaseThreshold, &keep));
    }
    c.max_grid = c.num_sms * 8 > kMaxBlocks ? c.num_sms * 8 : kMaxBlocks;
    NSB_CUDA(cudaMalloc((void**)&c.partials, sizeof(double) * kMaxRed * (size_t)c.max_grid));
    NSB_CUDA(cudaMalloc((void**)&c.chain, sizeof(double) * kMaxRed));
    NSB_CUDA(cudaMemset(c.chain, 0, sizeof(double) * kMaxRed));
    NSB_CUDA(cudaMalloc((void**)&c.ticket, sizeof(unsigned int)));
    NSB_CUDA(cudaMemset(c.ticket, 0, sizeof(unsigned int)));
    NSB_CUDA(cudaMallocHost((void**)&c.h_mail, sizeof(double) * 64));
    NSB_CUDA(cudaMalloc((void**)&c.d_mail, sizeof(double) * 64));
    NSB_CUDA(cudaMemset(c.d_mail, 0, sizeof(double) * 64));
    NSB_CUDA(cudaMalloc((void**)&c.fault, sizeof(int)));
    NSB_CUDA(cudaMemset(c.fault, 0, sizeof(int)));
#ifdef NSB_SYNC_STATS
    NSB_CUDA(cudaMalloc((void**)&c.sync_stats, sizeof(unsigned long long) * 16));
    NSB_CUDA(cudaMemset(c.sync_stats, 0, sizeof(unsigned long long) * 16));
#endif
    return NSB_OK;
}

int check_fault(Context& c, const char* who) {
    if (!c.fault || c.nranks <= 1) return NSB_OK;
    int f = 0;
    NSB_CUDA(cudaMemcpyAsync(&f, c.fault, sizeof(int), cudaMemcpyDeviceToHost, c.stream));
    NSB_CUDA(cudaStreamSynchronize(c.stream));
    if (!f) return NSB_OK;
    NSB_CUDA(cudaMemsetAsync(c.fault, 0, sizeof(int), c.stream));
    static const char* what[] = {"", "scalar all-reduce: a peer's partial sums", "halo push: a neighbour's \"consumed\" word",
                                 "SpMV: a neighbour's interface entries"};
    set_error("%s: rank %d waited more than %.1f s for %s (peer exchange timed out; a rank died or left the solve)", who, c.rank,
              1e-9 * (double)c.spin_timeout_ns, what[f >= 1 && f <= 3 ? f : 0]);
    return NSB_ERR_NCCL;
}

int dist_init(Context& c, int rank, int nranks, const char* uid);  // dist.cu
void dist_destroy(Context& c);

}  // namespace nsb

using namespace nsb;

extern "C" {

int nsb_version(void) { return 100; }

const char* nsb_last_error(void) { return g_err; }

int nsb_ctx_create(int device, nsb_ctx* out) {
    NSB_REQUIRE(out != nullptr, "nsb_ctx_create: out is NULL");
    int count = 0;
    cudaError_t e = cudaGetDeviceCount(&count);
    if (e != cudaSuccess || count == 0) {
        cudaGetLastError();
        set_error("nsb_ctx_create: no CUDA device available (%s); there is no CPU fallback", cudaGetErrorString(e));
        return NSB_ERR_CUDA;
    }
    NSB_REQUIRE(device >= 0 && device < count, "nsb_ctx_create: device %d out of range [0,%d)", device, count);
    nsb_ctx_s* h = new nsb_ctx_s();
    int r = ctx_init(h->c, device);
    if (r != NSB_OK) {
        delete h;
        return r;
    }
    *out = h;
    return NSB_OK;
}

int nsb_ctx_create_dist(int device, int rank, int nranks, const char unique_id[128], nsb_ctx* out) {
    NSB_REQUIRE(nranks >= 1 && rank >= 0 && rank < nranks, "nsb_ctx_create_dist: bad rank %d of %d", rank, nranks);
    NSB_CHECK(nsb_ctx_create(device, out));
    if (nranks > 1) {
        int r = dist_init((*out)->c, rank, nranks, unique_id);
        if (r != NSB_OK) {
            nsb_ctx_destroy(*out);
            *out = nullptr;
            return r;
        }
        // CUDA graphs stay on when both exchange steps run over peer memory (no host-side communication call
        // inside the iteration); with the NCCL fallback the iteration is launched kernel by kernel
        Context& cc = (*out)->c;
        const char* g = getenv("NSB_GRAPHS");
        cc.use_graphs = (g ? atoi(g) != 0 : true) && cc.p2p != nullptr;
        const char* sk = getenv("NSB_FAULT_SKIP_PUBLISH");
        cc.fault_skip_publish = (sk && atoi(sk) == rank) ? 1 : 0;
    }
    return NSB_OK;
}

int nsb_ctx_destroy(nsb_ctx ctx) {
    if (!ctx) return NSB_OK;
    Context& c = ctx->c;
    cudaSetDevice(c.device);
    if (c.stream) cudaStreamSynchronize(c.stream);
    dist_destroy(c);
    if (c.partials) cudaFree(c.partials);
    if (c.ticket) cudaFree(c.ticket);
    if (c.chain) cudaFree(c.chain);
    if (c.d_mail) cudaFree(c.d_mail);
    if (c.sync_stats) cudaFree(c.sync_stats);
    if (c.fault) cudaFree(c.fault);
    if (c.h_mail) cudaFreeHost(c.h_mail);
    if (c.stream) cudaStreamDestroy(c.stream);
    delete ctx;
    return NSB_OK;
}

int nsb_ctx_sync(nsb_ctx ctx) {
    NSB_REQUIRE(ctx, "nsb_ctx_sync: NULL context");
    NSB_CUDA(cudaStreamSynchronize(ctx->c.stream));
    return NSB_OK;
}

void* nsb_ctx_stream(nsb_ctx ctx) { return ctx ? (void*)ctx->c.stream : nullptr; }

int64_t nsb_ctx_launch_count(nsb_ctx ctx) { return ctx ? ctx->c.launches : 0; }

int nsb_ctx_sync_stats(nsb_ctx ctx, uint64_t out[16], int reset) {
    NSB_REQUIRE(ctx && out, "nsb_ctx_sync_stats: NULL argument");
    memset(out, 0, sizeof(uint64_t) * 16);
    Context& c = ctx->c;
    if (!c.sync_stats) return NSB_OK;   // library built without -DNSB_SYNC_STATS
    NSB_CUDA(cudaSetDevice(c.device));
    NSB_CUDA(cudaStreamSynchronize(c.stream));
    NSB_CUDA(cudaMemcpy(out, c.sync_stats, sizeof(uint64_t) * 16, cudaMemcpyDeviceToHost));
    if (reset) NSB_CUDA(cudaMemset(c.sync_stats, 0, sizeof(uint64_t) * 16));
    return NSB_OK;
}

int nsb_ctx_set_graphs(nsb_ctx ctx, int enabled) {
    NSB_REQUIRE(ctx, "nsb_ctx_set_graphs: NULL context");
    ctx->c.use_graphs = enabled != 0 && (ctx->c.nranks == 1 || ctx->c.p2p != nullptr);
    return NSB_OK;
}

}  // extern "C"
