// p2p.cu -- the two exchange steps of the row-partitioned solvers over NVLink peer memory instead of NCCL
// (the reference is single-process and has no counterpart).
//
// One process per GPU; buffers are cudaMalloc'ed, exported with cudaIpcGetMemHandle, the 64-byte handles are
// all-gathered once over the NCCL communicator and opened on every rank, after which the data path is plain
// stores into the peer's HBM through NVSwitch and flag words polled in local memory:
//
//   halo exchange   k_halo_push      gathers the owned interface entries and stores them straight into each
//                                    slab neighbour's landing buffer, then releases a sequence flag there;
//                   k_halo_wait_copy polls the flags of the neighbours and moves the landing buffer into the
//                                    halo tail of x.  Two small kernels, no host involvement, graph-capturable.
//   scalar sum      k_p2p_allreduce  one block: every rank stores its partial sums into slot [rank] of every
//                                    peer's mailbox, polls for all slots and adds them in rank order, so all
//                                    ranks get the same bits (deterministic, partition-independent order).
//
// Flow control.  Flags carry monotonically increasing sequence numbers kept in device memory (so a captured
// graph replays correctly).  A landing buffer is single-buffered: before pushing exchange e a rank tells its
// neighbours "I consumed e-1" (true by stream order) and waits for the same word from them.  Mailboxes are
// double-buffered by the parity of the sequence number: a rank can be at most one reduction ahead of a peer,
// because finishing a reduction needs every peer's contribution to it.
#include <stdlib.h>

#include <vector>

#include "p2p.cuh"

namespace nsb {

// ---- IPC helper: all ranks export one allocation and open everyone else's -------------------------------
struct IpcBlock {
    void* local = nullptr;
    size_t bytes = 0;
    std::vector<void*> peer;   // peer[r]: rank r's block mapped here (peer[rank] == local)
    int rank = 0;
};

static void ipc_release(IpcBlock& b) {
    for (size_t r = 0; r < b.peer.size(); ++r)
        if (b.peer[r] && (int)r != b.rank) cudaIpcCloseMemHandle(b.peer[r]);
    b.peer.clear();
    if (b.local) cudaFree(b.local);
    b.local = nullptr;
}

// Collective.  `want[r]` says whether this rank needs rank r's block.  Returns NSB_OK with ok = false when
// IPC is not available (caller falls back to NCCL); every rank learns the same verdict.
static int ipc_exchange(Context& c, size_t bytes, const std::vector<char>& want, IpcBlock& b, bool* ok) {
    *ok = false;
    b.rank = c.rank;
    b.bytes = bytes;
    struct Rec {
        cudaIpcMemHandle_t h;
        int good;
        int pad[3];
    };
    Rec mine;
    memset(&mine, 0, sizeof(mine));
    cudaError_t e = cudaMalloc(&b.local, bytes);
    if (e == cudaSuccess) e = cudaMemset(b.local, 0, bytes);
    if (e == cudaSuccess) e = cudaIpcGetMemHandle(&mine.h, b.local);
    mine.good = e == cudaSuccess ? 1 : 0;
    if (e != cudaSuccess) cudaGetLastError();
    DevBuf<Rec> dsend, drecv;
    NSB_CHECK(dsend.alloc(1));
    NSB_CHECK(drecv.alloc(c.nranks));
    NSB_CUDA(cudaMemcpy(dsend.p, &mine, sizeof(Rec), cudaMemcpyHostToDevice));
    NSB_CHECK(nccl_allgather_bytes(c, dsend.p, drecv.p, sizeof(Rec)));
    std::vector<Rec> all(c.nranks);
    NSB_CUDA(cudaMemcpy(all.data(), drecv.p, sizeof(Rec) * c.nranks, cudaMemcpyDeviceToHost));
    int good = 1;
    for (int r = 0; r < c.nranks; ++r) good &= all[r].good;
    b.peer.assign(c.nranks, nullptr);
    if (good) {
        b.peer[c.rank] = b.local;
        for (int r = 0; r < c.nranks && good; ++r) {
            if (r == c.rank || !want[r]) continue;
            e = cudaIpcOpenMemHandle(&b.peer[r], all[r].h, cudaIpcMemLazyEnablePeerAccess);
            if (e != cudaSuccess) {
                cudaGetLastError();
                b.peer[r] = nullptr;
                good = 0;
            }
        }
    }
    // agree on the verdict (an open can fail on one rank only)
    int verdict = good;
    DevBuf<int> vs, vr;
    NSB_CHECK(vs.alloc(1));
    NSB_CHECK(vr.alloc(c.nranks));
    NSB_CUDA(cudaMemcpy(vs.p, &verdict, sizeof(int), cudaMemcpyHostToDevice));
    NSB_CHECK(nccl_allgather_bytes(c, vs.p, vr.p, sizeof(int)));
    std::vector<int> vall(c.nranks);
    NSB_CUDA(cudaMemcpy(vall.data(), vr.p, sizeof(int) * c.nranks, cudaMemcpyDeviceToHost));
    for (int r = 0; r < c.nranks; ++r) verdict &= vall[r];
    if (!verdict) {
        ipc_release(b);
        return NSB_OK;
    }
    *ok = true;
    return NSB_OK;
}

// ---- scalar all-reduce -------------------------------------------------------------------------------

struct P2PReduce {
    IpcBlock blk;              // [ box: 2 x nranks x kArSlots doubles | flags: nranks u64 | seq u64 ]
    DevBuf<double*> peer_box;  // per rank: that rank's box
    DevBuf<u64*> peer_flag;    // per rank: that rank's flag array
    double* my_box = nullptr;
    u64* my_flag = nullptr;
    u64* seq = nullptr;
};

P2PReduceView p2p_reduce_view(Context& c) {
    P2PReduce* P = c.p2p;
    return P2PReduceView{P->peer_box.p, P->peer_flag.p, P->my_box, P->my_flag, P->seq, c.rank, c.nranks, c.sync_stats,
                         SpinGuard{c.fault, c.spin_timeout_ns}};
}

RedMode red_mode(Context& c) {
    RedMode M;
    memset(&M, 0, sizeof(M));
    static int inblock = -1;
    if (inblock < 0) {
        const char* e = getenv("NSB_REDUCE_INBLOCK");   // 0: the all-reduce runs as a kernel of its own (A/B measurements)
        inblock = (e && atoi(e) == 0) ? 0 : 1;
    }
    M.kind = c.nranks <= 1 ? kRedLocal : ((c.p2p && inblock) ? kRedPeer : kRedStore);
    if (M.kind == kRedPeer) M.R = p2p_reduce_view(c);
    return M;
}

int p2p_reduce_init(Context& c) {
    c.p2p = nullptr;
    if (c.nranks <= 1 || c.nranks > 64) return NSB_OK;
    P2PReduce* P = new P2PReduce();
    const size_t box_bytes = sizeof(double) * 2 * c.nranks * kArSlots;
    const size_t bytes = box_bytes + sizeof(u64) * (c.nranks + 1);
    std::vector<char> want(c.nranks, 1);
    bool ok = false;
    int rc = ipc_exchange(c, bytes, want, P->blk, &ok);
    if (rc != NSB_OK || !ok) {
        delete P;
        return rc;
    }
    std::vector<double*> pb(c.nranks);
    std::vector<u64*> pf(c.nranks);
    for (int r = 0; r < c.nranks; ++r) {
        pb[r] = (double*)P->blk.peer[r];
        pf[r] = (u64*)((char*)P->blk.peer[r] + box_bytes);
    }
    P->my_box = pb[c.rank];
    P->my_flag = pf[c.rank];
    P->seq = P->my_flag + c.nranks;
    rc = P->peer_box.alloc(c.nranks);
    if (rc == NSB_OK) rc = P->peer_flag.alloc(c.nranks);
    if (rc == NSB_OK) {
        cudaMemcpy(P->peer_box.p, pb.data(), sizeof(double*) * c.nranks, cudaMemcpyHostToDevice);
        cudaMemcpy(P->peer_flag.p, pf.data(), sizeof(u64*) * c.nranks, cudaMemcpyHostToDevice);
    }
    if (rc != NSB_OK) {
        ipc_release(P->blk);
        delete P;
        return rc;
    }
    c.p2p = P;   // (whether the iteration loops are captured in graphs is decided in nsb_ctx_create_dist: NSB_GRAPHS is honoured)
    return NSB_OK;
}

void p2p_reduce_destroy(Context& c) {
    if (!c.p2p) return;
    ipc_release(c.p2p->blk);
    delete c.p2p;
    c.p2p = nullptr;
}

int p2p_allreduce(Context& c, double* dev, int count) {
    p2p_allreduce_then(c, dev, count, NoPost{});
    NSB_CUDA(cudaGetLastError());
    return NSB_OK;
}

// ---- halo exchange -----------------------------------------------------------------------------------
struct P2PHalo {
    IpcBlock blk;   // [ landing: 2 x n_halo doubles (padded; one area per vector of a dual exchange) | ready: nnb u64 | consumed: nnb u64 | seq u64 | ticket u32 ]
    DevBuf<int64_t> send_off;
    DevBuf<double*> peer_land;
    DevBuf<u64*> peer_ready, peer_consumed;
    DevBuf<int64_t> peer_land2;
    HaloDev dev;
    int push_grid = 1, copy_grid = 1;
    // the send lists as runs of consecutive owned rows (fused push from the producing kernel); nruns == 0: not run-shaped
    int nruns = 0;
    int run_start[kPushRuns], run_len[kPushRuns], run_nb[kPushRuns], run_slot0[kPushRuns];
    DevBuf<unsigned char> tiles;   // one byte per 256 owned rows: the tile holds a row that is sent
};

__global__ void __launch_bounds__(kBlock) k_halo_push(HaloDev H, const double* __restrict__ x, const double* __restrict__ x2) {
    const u64 e = *H.seq + 1;
    // my previous landing buffer content has been moved out (stream order): tell the neighbours, then wait
    // until they have moved out what I pushed last time
    if (blockIdx.x == 0 && threadIdx.x < H.nnb) st_release_sys(H.peer_consumed[threadIdx.x], e - 1);
    if (threadIdx.x < H.nnb) {
        NSB_STATS_T0();
        spin_until(H.my_consumed + threadIdx.x, e - 1, H.guard, kFaultHaloConsumed);
        if (blockIdx.x == 0 && threadIdx.x == 0) NSB_STATS_ADD(H.stats, 2);
    }
    __syncthreads();
    for (int64_t i = (int64_t)blockIdx.x * blockDim.x + threadIdx.x; i < H.n_send; i += (int64_t)gridDim.x * blockDim.x) {
        int k = 0;
        while (k + 1 < H.nnb && i >= H.send_off[k + 1]) ++k;
        const int src = H.send_idx[i];
        H.peer_land[k][i - H.send_off[k]] = x[src];
        if (x2) H.peer_land[k][H.peer_land2[k] + i - H.send_off[k]] = x2[src];
    }
    __threadfence_system();
    __syncthreads();
    __shared__ bool last;
    if (threadIdx.x == 0) last = atomicAdd(H.ticket, 1u) == gridDim.x - 1;
    __syncthreads();
    if (last) {
        __threadfence_system();
        if (threadIdx.x < H.nnb && !H.skip_publish) st_release_sys(H.peer_ready[threadIdx.x], e);
        if (threadIdx.x == 0) {
            *H.ticket = 0u;
            *H.seq = e;
        }
    }
}

__global__ void __launch_bounds__(kBlock) k_halo_wait_copy(HaloDev H, double* __restrict__ tail, double* __restrict__ tail2) {
    const u64 e = *H.seq;   // the push of this exchange is earlier in the stream
    if (threadIdx.x < H.nnb) {
        NSB_STATS_T0();
        spin_until(H.my_ready + threadIdx.x, e, H.guard, kFaultHaloReady);
        if (blockIdx.x == 0 && threadIdx.x == 0) NSB_STATS_ADD(H.stats, 6);
    }
    __syncthreads();
    for (int64_t i = (int64_t)blockIdx.x * blockDim.x + threadIdx.x; i < H.n_halo; i += (int64_t)gridDim.x * blockDim.x)
    {
        tail[i] = __ldcv(H.my_land + i);
        if (tail2) tail2[i] = __ldcv(H.my_land + H.land2 + i);
    }
}

int p2p_halo_init(Context& c, nsb_halo_s* h) {
    h->p2p = nullptr;
    const int nnb = (int)h->neighbors.size();
    const int R = c.nranks;
    const int64_t land = (h->n_halo + 7) & ~7LL;
    // table shared with everyone: [land, nnb, recv_off[R], slot[R]]
    std::vector<int64_t> mine(2 + 2 * (size_t)R, -1);
    mine[0] = land;
    mine[1] = nnb;
    // Arguments are validated BEFORE the collective, but the verdict travels with it: a rank that returned early here would
    // leave its peers blocked inside the all-gather.  mine[0] < 0 tells every rank that this one was handed a bad plan.
    int64_t ro = 0;
    int bad_nb = -1;
    for (int k = 0; k < nnb; ++k) {
        const int nb = h->neighbors[k];
        if (nb < 0 || nb >= R || nb == c.rank) {
            bad_nb = nb;
            mine[0] = -1;
            break;
        }
        mine[2 + nb] = ro;
        mine[2 + R + nb] = k;
        ro += h->recv_counts[k];
    }
    if (nnb > kBlock) mine[0] = -1;
    DevBuf<int64_t> ts, tr;
    NSB_CHECK(ts.alloc((int64_t)mine.size()));
    NSB_CHECK(tr.alloc((int64_t)mine.size() * R));
    NSB_CUDA(cudaMemcpy(ts.p, mine.data(), sizeof(int64_t) * mine.size(), cudaMemcpyHostToDevice));
    NSB_CHECK(nccl_allgather_bytes(c, ts.p, tr.p, sizeof(int64_t) * mine.size()));
    std::vector<int64_t> all(mine.size() * R);
    NSB_CUDA(cudaMemcpy(all.data(), tr.p, sizeof(int64_t) * all.size(), cudaMemcpyDeviceToHost));
    auto tab = [&](int r) { return all.data() + (size_t)r * mine.size(); };
    for (int r = 0; r < R; ++r)
        if (tab(r)[0] < 0) {   // every rank sees the same table: every rank returns the error
            if (r == c.rank && bad_nb != -1) set_error("nsb_halo_create: neighbour %d out of range (rank %d of %d)", bad_nb, c.rank, R);
            else if (r == c.rank) set_error("nsb_halo_create: more than %d neighbours", kBlock);
            else set_error("nsb_halo_create: rank %d was given an invalid halo plan", r);
            return NSB_ERR_INVALID;
        }
    for (int r = 0; r < R; ++r)          // symmetry of the neighbour lists, judged from the gathered tables: the same verdict on every rank
        for (int q = 0; q < R; ++q)
            if (tab(r)[2 + R + q] >= 0 && tab(q)[2 + R + r] < 0) {
                set_error("nsb_halo_create: neighbour lists are not symmetric across ranks (rank %d lists %d, not the reverse)", r, q);
                return NSB_ERR_INVALID;
            }

    P2PHalo* P = new P2PHalo();
    const size_t bytes = sizeof(double) * 2 * (size_t)land + sizeof(u64) * (2 * (size_t)nnb + 1) + 16;   // two vector areas
    std::vector<char> want(R, 0);
    for (int k = 0; k < nnb; ++k) want[h->neighbors[k]] = 1;
    bool ok = false;
    int rc = ipc_exchange(c, bytes, want, P->blk, &ok);
    if (rc != NSB_OK || !ok) {
        delete P;
        return rc;
    }
    std::vector<double*> pl(nnb > 0 ? nnb : 1, nullptr);
    std::vector<u64*> pr(nnb > 0 ? nnb : 1, nullptr), pc(nnb > 0 ? nnb : 1, nullptr);
    std::vector<int64_t> so(nnb + 1, 0), l2(nnb > 0 ? nnb : 1, 0);
    bool sym = true;
    for (int k = 0; k < nnb; ++k) {
        const int nb = h->neighbors[k];
        const int64_t* T = tab(nb);
        const int64_t off = T[2 + c.rank], slot = T[2 + R + c.rank];
        if (off < 0 || slot < 0) {
            sym = false;
            break;
        }
        char* base = (char*)P->blk.peer[nb];
        pl[k] = (double*)base + off;
        u64* flags = (u64*)(base + sizeof(double) * 2 * (size_t)T[0]);
        l2[k] = T[0];
        pr[k] = flags + slot;
        pc[k] = flags + T[1] + slot;
        so[k + 1] = so[k] + h->send_counts[k];
    }
    if (!sym) {
        ipc_release(P->blk);
        delete P;
        set_error("nsb_halo_create: neighbour lists are not symmetric across ranks");
        return NSB_ERR_INVALID;
    }
    rc = P->send_off.alloc(nnb + 1);
    if (rc == NSB_OK) rc = P->peer_land.alloc(nnb > 0 ? nnb : 1);
    if (rc == NSB_OK) rc = P->peer_ready.alloc(nnb > 0 ? nnb : 1);
    if (rc == NSB_OK) rc = P->peer_consumed.alloc(nnb > 0 ? nnb : 1);
    if (rc == NSB_OK) rc = P->peer_land2.alloc(nnb > 0 ? nnb : 1);
    if (rc != NSB_OK) {
        ipc_release(P->blk);
        delete P;
        return rc;
    }
    cudaMemcpy(P->send_off.p, so.data(), sizeof(int64_t) * (nnb + 1), cudaMemcpyHostToDevice);
    cudaMemcpy(P->peer_land.p, pl.data(), sizeof(double*) * pl.size(), cudaMemcpyHostToDevice);
    cudaMemcpy(P->peer_ready.p, pr.data(), sizeof(u64*) * pr.size(), cudaMemcpyHostToDevice);
    cudaMemcpy(P->peer_consumed.p, pc.data(), sizeof(u64*) * pc.size(), cudaMemcpyHostToDevice);
    cudaMemcpy(P->peer_land2.p, l2.data(), sizeof(int64_t) * l2.size(), cudaMemcpyHostToDevice);
    char* lb = (char*)P->blk.local;
    u64* flags = (u64*)(lb + sizeof(double) * 2 * (size_t)land);
    HaloDev& D = P->dev;
    D.nnb = nnb;
    D.n_send = h->n_send;
    D.n_halo = h->n_halo;
    D.send_idx = h->send_idx.p;
    D.send_off = P->send_off.p;
    D.peer_land = P->peer_land.p;
    D.peer_ready = P->peer_ready.p;
    D.peer_consumed = P->peer_consumed.p;
    D.my_land = (const double*)lb;
    D.land2 = land;
    D.peer_land2 = P->peer_land2.p;
    D.my_ready = flags;
    D.my_consumed = flags + nnb;
    D.seq = flags + 2 * nnb;
    D.ticket = (unsigned int*)(flags + 2 * nnb + 1);
    D.stats = c.sync_stats;
    D.guard = SpinGuard{c.fault, c.spin_timeout_ns};
    D.skip_publish = c.fault_skip_publish;
    // the send lists as runs of consecutive owned rows
    {
        std::vector<int32_t> sidx((size_t)(h->n_send > 0 ? h->n_send : 0));
        if (h->n_send > 0) cudaMemcpy(sidx.data(), h->send_idx.p, sizeof(int32_t) * (size_t)h->n_send, cudaMemcpyDeviceToHost);
        int nr = 0;
        bool ok = true;
        for (int k = 0; k < nnb && ok; ++k) {
            const int64_t a = so[k], b = so[k + 1];
            for (int64_t i = a; i < b && ok;) {
                int64_t j = i + 1;
                while (j < b && sidx[(size_t)j] == sidx[(size_t)j - 1] + 1) ++j;
                if (nr == kPushRuns) {
                    ok = false;
                    break;
                }
                P->run_start[nr] = sidx[(size_t)i];
                P->run_len[nr] = (int)(j - i);
                P->run_nb[nr] = k;
                P->run_slot0[nr] = (int)(i - a);
                ++nr;
                i = j;
            }
        }
        P->nruns = ok ? nr : 0;
        if (P->nruns > 0) {
            const int64_t nt = (h->n_owned >> kPushTileLog2) + 2;
            std::vector<unsigned char> tf((size_t)nt, 0);
            for (int r = 0; r < nr; ++r)
                for (int64_t t = (int64_t)P->run_start[r] >> kPushTileLog2; t <= ((int64_t)P->run_start[r] + P->run_len[r] - 1) >> kPushTileLog2; ++t)
                    tf[(size_t)t] = 1;
            if (P->tiles.alloc(nt) != NSB_OK || cudaMemcpy(P->tiles.p, tf.data(), (size_t)nt, cudaMemcpyHostToDevice) != cudaSuccess) P->nruns = 0;
        }
    }
    auto grid_for = [&](int64_t n) {
        int64_t g = (n + (int64_t)kBlock * 4 - 1) / ((int64_t)kBlock * 4);
        if (g < 1) g = 1;
        if (g > 64) g = 64;   // small grids: every block polls the flags
        return (int)g;
    };
    P->push_grid = grid_for(h->n_send);
    P->copy_grid = grid_for(h->n_halo);
    if (nnb > kBlock) {
        ipc_release(P->blk);
        delete P;
        set_error("nsb_halo_create: more than %d neighbours", kBlock);
        return NSB_ERR_UNSUPPORTED;
    }
    h->p2p = P;
    return NSB_OK;
}

void p2p_halo_destroy(nsb_halo_s* h) {
    if (!h || !h->p2p) return;
    ipc_release(h->p2p->blk);
    delete h->p2p;
    h->p2p = nullptr;
}

int p2p_halo_push(Context& c, nsb_halo_s* h, double* x, double* x2) {
    P2PHalo* P = h->p2p;
    if (P->dev.nnb == 0) return NSB_OK;
    k_halo_push<<<P->push_grid, kBlock, 0, c.stream>>>(P->dev, x, x2);
    c.launches++;
    NSB_CUDA(cudaGetLastError());
    return NSB_OK;
}

// View for a SpMV that reads the halo entries of vector `area` (0: x, 1: x2 of the last push) from the landing buffer.
HaloIn p2p_halo_in(const nsb_halo_s* h, int area, const int* order, int n_interior) {
    const HaloDev& D = h->p2p->dev;
    return HaloIn{D.my_ready, D.seq, D.my_land + (area ? D.land2 : 0), h->n_owned, D.nnb, order, n_interior, D.stats, D.guard};
}

PushOut p2p_push_out(const nsb_halo_s* h) {
    PushOut P;
    memset(&P, 0, sizeof(P));
    if (!h || !h->p2p || h->p2p->dev.nnb == 0) return P;
    const P2PHalo* Q = h->p2p;
    P.H = Q->dev;
    P.nruns = Q->nruns;
    P.tiles = Q->tiles.p;
    for (int k = 0; k < Q->nruns; ++k) {
        P.start[k] = Q->run_start[k];
        P.len[k] = Q->run_len[k];
        P.nb[k] = Q->run_nb[k];
        P.slot0[k] = Q->run_slot0[k];
    }
    return P;
}

// The vector this operator multiplies can be pushed from the epilogue of the kernel that produces it: the operator takes its halo
// inside the SpMV kernel (so nothing else needs the exchange) and the send lists are a few runs of consecutive rows.
bool push_fused_ok(const Context& c, const nsb_mat_s* m) {
    static int enabled = -1;
    if (enabled < 0) {
        const char* e = getenv("NSB_PUSH_FUSED");
        enabled = (e && atoi(e) == 0) ? 0 : 1;
    }
    return enabled && halo_fused_ok(c, m) && m->halo->p2p->nruns > 0;
}

bool halo_fused_ok(const Context& c, const nsb_mat_s* m) {
    static int enabled = -1;
    if (enabled < 0) {
        const char* e = getenv("NSB_HALO_FUSED");
        enabled = (e && atoi(e) == 0) ? 0 : 1;
    }
    if (!enabled || c.nranks <= 1 || !m->halo || !m->halo->p2p || m->panels.empty()) return false;
    for (const auto& p : m->panels)
        if (p->halo_n_interior < 0) return false;
    return true;
}

int p2p_halo_exchange(Context& c, nsb_halo_s* h, double* x, double* x2) {
    P2PHalo* P = h->p2p;
    if (P->dev.nnb == 0) return NSB_OK;
    k_halo_push<<<P->push_grid, kBlock, 0, c.stream>>>(P->dev, x, x2);
    k_halo_wait_copy<<<P->copy_grid, kBlock, 0, c.stream>>>(P->dev, x + h->n_owned, x2 ? x2 + h->n_owned : nullptr);
    c.launches += 2;
    NSB_CUDA(cudaGetLastError());
    return NSB_OK;
}

}  // namespace nsb
