// sell.cu -- builds the SELL-LX companion of a CSR operator on the device (kernel: spmv.cuh).
//
// Why this format.  ncu on the CSR-vector kernel (profiles/r1_spmv_csrvec_A_2Mdofs_ncu.csv) showed DRAM
// traffic equal to the algorithmic bytes but the L1 data pipe at 72 %: every warp-wide gather of x touches
// 5+ cache lines, so L1, not HBM, bounded it at ~60 % of the copy roofline.  The first fix (x staged in
// shared memory, 16-bit chunk-local columns over plain CSR rows) reached 67 % of DRAM peak and was latency
// bound (profiles/r1_bpcg_iteration_ncu.md: long-scoreboard stalls, dependent row-pointer -> value loads,
// 64-bit / 16-bit loads).  SELL-LX keeps the staging and stores the nonzeros in the order the lanes
// consume them:
//   * rows are cut into chunks of R consecutive rows; per chunk the sorted set of distinct columns is
//     stored once (ucol, 4 B each) and every nonzero stores a 16-bit index into that set (lcol);
//   * inside a chunk the rows are sorted by length (sigma = R) and cut into slices of 32/T rows, T lanes
//     per row; a slice is padded to its longest row and stored as pair-blocks of 32 lanes x 2 entries,
//     so loads are 128-bit (values) / 32-bit (indices), fully coalesced, with no row pointers.
// Traffic per SpMV: 10 B per stored nonzero (padding is 2-8 % on the P2/P1 blocks) + 8 B per run of consecutive
// distinct columns (or 4 B per distinct column when the runs are short) + 2 B per row (perm) + x once + y once.
// R is the largest power of two whose staging buffers (distinct columns + R row sums) fit kSellStage
// doubles: 4 resident blocks then use ~144 KB of shared memory and leave ~84 KB of L1 for the x gathers of
// the staging phase (with 228 KB of shared memory in use the same kernel ran 20 % slower).  An operator
// that does not fit even at the smallest R keeps the CSR-vector kernel.
#include <cub/cub.cuh>
#include <stdlib.h>

#include "spmv.cuh"

namespace nsb {

// ---- chunk-local column numbering ----------------------------------------------------------------
__global__ void k_sell_keys(CsrView A, int pbase, int log2R, long long* __restrict__ keys, int* __restrict__ pos) {
    const int lane = threadIdx.x & 31;
    const int64_t warp = ((int64_t)blockIdx.x * blockDim.x + threadIdx.x) >> 5;
    const int64_t nwarps = ((int64_t)gridDim.x * blockDim.x) >> 5;
    for (int64_t r = warp; r < A.rows; r += nwarps) {
        const int p0 = A.rowptr[r], p1 = A.rowptr[r + 1];
        const long long hi = (long long)(r >> log2R) << 32;
        for (int p = p0 + lane; p < p1; p += 32) {
            keys[p - pbase] = hi | (long long)A.colidx[p];
            pos[p - pbase] = p - pbase;
        }
    }
}

__global__ void k_sell_heads(const long long* __restrict__ keys, int* __restrict__ head, int64_t nnz) {
    for (int64_t i = (int64_t)blockIdx.x * blockDim.x + threadIdx.x; i < nnz; i += (int64_t)gridDim.x * blockDim.x)
        head[i] = (i == 0 || keys[i] != keys[i - 1]) ? 1 : 0;
}

// uptr[c] = number of distinct (chunk, col) pairs with chunk < c
__global__ void k_sell_uptr(const long long* __restrict__ keys, const int* __restrict__ gidx_incl, int64_t nnz,
                            int* __restrict__ uptr, int nchunks) {
    for (int c = blockIdx.x * blockDim.x + threadIdx.x; c <= nchunks; c += gridDim.x * blockDim.x) {
        const long long target = (long long)c << 32;
        int64_t lo = 0, hi = nnz;  // first i with keys[i] >= target
        while (lo < hi) {
            const int64_t mid = (lo + hi) >> 1;
            if (keys[mid] < target) lo = mid + 1;
            else hi = mid;
        }
        uptr[c] = lo == 0 ? 0 : gidx_incl[lo - 1];
    }
}

// lcol in CSR order (temporary) and the distinct-column lists
__global__ void k_sell_localcols(const long long* __restrict__ keys, const int* __restrict__ pos,
                                 const int* __restrict__ head, const int* __restrict__ gidx_incl,
                                 const int* __restrict__ uptr, unsigned short* __restrict__ lcol_csr,
                                 int* __restrict__ ucol, int64_t nnz) {
    for (int64_t i = (int64_t)blockIdx.x * blockDim.x + threadIdx.x; i < nnz; i += (int64_t)gridDim.x * blockDim.x) {
        const long long k = keys[i];
        const int chunk = (int)(k >> 32);
        const int g = gidx_incl[i] - 1;
        lcol_csr[pos[i]] = (unsigned short)(g - uptr[chunk]);
        if (head[i]) ucol[g] = (int)(k & 0xffffffffLL);
    }
}

__global__ void k_sell_maxuniq(const int* __restrict__ uptr, int nchunks, int* __restrict__ out) {
    int m = 0;
    for (int c = blockIdx.x * blockDim.x + threadIdx.x; c < nchunks; c += gridDim.x * blockDim.x)
        m = max(m, uptr[c + 1] - uptr[c]);
    atomicMax(out, m);
}

// ---- runs of consecutive columns in the distinct-column lists -------------------------------------------
// FEM numberings make the distinct columns of a chunk a few long runs (whole mesh lines); storing
// (first column, position) per run instead of 4 B per column removes most of the index traffic of the
// staging phase and turns its gathers into contiguous copies.
__global__ void k_sell_chunkflags(const int* __restrict__ uptr, int nchunks, int total, int* __restrict__ flag) {
    for (int c = blockIdx.x * blockDim.x + threadIdx.x; c < nchunks; c += gridDim.x * blockDim.x)
        if (uptr[c] < total) flag[uptr[c]] = 1;
}

__global__ void k_sell_runheads(const int* __restrict__ ucol, int total, int* __restrict__ flag) {
    for (int g = blockIdx.x * blockDim.x + threadIdx.x; g < total; g += gridDim.x * blockDim.x)
        if (g == 0 || ucol[g] != ucol[g - 1] + 1) flag[g] = 1;   // benign race: every writer stores 1
}

__global__ void k_sell_runfill(const int* __restrict__ ucol, const int* __restrict__ flag, const int* __restrict__ ridx_incl,
                               int total, int* __restrict__ rstart, int* __restrict__ rpos) {
    for (int g = blockIdx.x * blockDim.x + threadIdx.x; g < total; g += gridDim.x * blockDim.x)
        if (flag[g]) {
            rstart[ridx_incl[g] - 1] = ucol[g];
            rpos[ridx_incl[g] - 1] = g;
        }
}

__global__ void k_sell_runptr(const int* __restrict__ uptr, const int* __restrict__ ridx_incl, int nchunks, int total, int nruns,
                              int* __restrict__ rptr, int* __restrict__ rpos) {
    for (int c = blockIdx.x * blockDim.x + threadIdx.x; c <= nchunks; c += gridDim.x * blockDim.x) {
        rptr[c] = (c < nchunks && uptr[c] < total) ? ridx_incl[uptr[c]] - 1 : nruns;
        if (c == nchunks) rpos[nruns] = total;
    }
}

// ---- sorting the rows of every chunk by length ------------------------------------------------------
constexpr unsigned long long kLenMax = 0xFFFFFFull;

// key = chunk | (kLenMax - len) | row-in-chunk : ascending sort = chunks in order, longest row first,
// ties by row number (so neighbouring mesh rows stay neighbours).  Virtual rows past the end have len 0.
__global__ void k_sell_rowkeys(CsrView A, int log2R, int64_t nvr, unsigned long long* __restrict__ keys) {
    for (int64_t i = (int64_t)blockIdx.x * blockDim.x + threadIdx.x; i < nvr; i += (int64_t)gridDim.x * blockDim.x) {
        unsigned long long len = i < A.rows ? (unsigned long long)(A.rowptr[i + 1] - A.rowptr[i]) : 0ull;
        if (len > kLenMax) len = kLenMax;
        keys[i] = ((unsigned long long)(i >> log2R) << 40) | ((kLenMax - len) << 16) |
                  (unsigned long long)(i & ((1LL << log2R) - 1));
    }
}

// perm and the pair-block count of every slice (taken from the slice's first = longest row)
__global__ void k_sell_slices(const unsigned long long* __restrict__ keys, int64_t nvr, int rps, int T,
                              unsigned short* __restrict__ perm, unsigned int* __restrict__ npb) {
    for (int64_t i = (int64_t)blockIdx.x * blockDim.x + threadIdx.x; i < nvr; i += (int64_t)gridDim.x * blockDim.x) {
        const unsigned long long k = keys[i];
        perm[i] = (unsigned short)(k & 0xFFFFull);
        if (i % rps == 0) {
            const unsigned int len = (unsigned int)(kLenMax - ((k >> 16) & kLenMax));
            npb[i / rps] = (len + 2 * T - 1) / (2 * T);
        }
    }
}

// one warp per sorted row position: copy the row's entries to their lanes' slots
__global__ void k_sell_fill(CsrView A, int pbase, const unsigned short* __restrict__ lcol_csr, int log2R, int64_t nvr, int rps, int T,
                            const unsigned short* __restrict__ perm, const unsigned int* __restrict__ off,
                            double* __restrict__ vals, unsigned short* __restrict__ lcol) {
    const int lane = threadIdx.x & 31;
    const int64_t warp = ((int64_t)blockIdx.x * blockDim.x + threadIdx.x) >> 5;
    const int64_t nwarps = ((int64_t)gridDim.x * blockDim.x) >> 5;
    for (int64_t i = warp; i < nvr; i += nwarps) {
        const int64_t r = ((i >> log2R) << log2R) + perm[i];
        if (r >= A.rows) continue;
        const int p0 = A.rowptr[r], len = A.rowptr[r + 1] - p0;
        const int64_t base = (int64_t)off[i / rps] * 64;
        const int ris = (int)(i % rps);
        for (int e = lane; e < len; e += 32) {
            const int pair = e >> 1;
            const int64_t dst = base + (int64_t)(pair / T) * 64 + (ris * T + pair % T) * 2 + (e & 1);
            vals[dst] = A.values[p0 + e];
            lcol[dst] = lcol_csr[p0 + e - pbase];
        }
    }
}

// Bank-aware placement of a row's entries (T >= 4 lanes per row).  The compute phase gathers x from shared memory with one
// 64-bit load per stored entry; the 16 lanes of a half-warp -- 16 / T rows, T lanes each -- are served in one pass only when their
// 16 chunk-local column indices differ mod 16 (16 eight-byte banks).  With the entries of a row in ascending column order the
// long pressure rows of B (189 entries, 8 lanes per row) ran at 3 extra passes per load (ncu, round 1: 16.2M bank conflicts on
// 5.3M loads, short-scoreboard and MIO-throttle stalls, 0.76 of the DRAM rate of the other SpMVs).  The order of the entries inside a
// row is free, so each row is laid out group by group (a group = the T entries its lanes load together): the 16 residues are cut
// into 16 / T sectors of T residues, row rho of the half-warp takes its T entries of group g from sector (g + rho) mod (16 / T), one
// per residue -- the rows of a half-warp then hit disjoint banks -- and falls back to its fullest other residue when one runs dry.
constexpr int kBankRowMax = 1024;   // rows longer than this keep the ascending order

__global__ void __launch_bounds__(128) k_sell_fill_banked(CsrView A, int pbase, const unsigned short* __restrict__ lcol_csr, int log2R,
                                                          int64_t nvr, int rps, int T, const unsigned short* __restrict__ perm,
                                                          const unsigned int* __restrict__ off, double* __restrict__ vals,
                                                          unsigned short* __restrict__ lcol) {
    const int Q = 16 / T;   // rows per half-warp = residue sectors
    for (int64_t i = (int64_t)blockIdx.x * blockDim.x + threadIdx.x; i < nvr; i += (int64_t)gridDim.x * blockDim.x) {
        const int64_t r = ((i >> log2R) << log2R) + perm[i];
        if (r >= A.rows) continue;
        const int p0 = A.rowptr[r], len = A.rowptr[r + 1] - p0;
        const int64_t base = (int64_t)off[i / rps] * 64;
        const int ris = (int)(i % rps);
        auto put = [&](int pos, int e) {   // entry e of the row goes to position pos = (group, member): group = 2 * pair-block + slot
            const int g = pos / T, t = pos % T;
            const int64_t dst = base + (int64_t)(g >> 1) * 64 + (ris * T + t) * 2 + (g & 1);
            vals[dst] = A.values[p0 + e];
            lcol[dst] = lcol_csr[p0 + e - pbase];
        };
        if (len > kBankRowMax) {
            for (int e = 0; e < len; ++e) put((e >> 1) / T * 2 * T + (e & 1) * T + (e >> 1) % T, e);   // ascending order, as k_sell_fill
            continue;
        }
        unsigned short byres[kBankRowMax];   // the row's entries bucketed by residue (counting sort, stable)
        int cnt[16], head[16];
        for (int b = 0; b < 16; ++b) cnt[b] = 0;
        for (int e = 0; e < len; ++e) cnt[lcol_csr[p0 + e - pbase] & 15]++;
        int acc = 0;
        for (int b = 0; b < 16; ++b) {
            head[b] = acc;
            acc += cnt[b];
        }
        {
            int fill[16];
            for (int b = 0; b < 16; ++b) fill[b] = head[b];
            for (int e = 0; e < len; ++e) byres[fill[lcol_csr[p0 + e - pbase] & 15]++] = (unsigned short)e;
        }
        const int rho = ris % Q;
        int placed = 0;
        for (int g = 0; placed < len; ++g) {
            const int sector = (g + rho) % Q;
            unsigned int used = 0;   // residues this row already used in this group
            for (int t = 0; t < T && placed < len; ++t) {
                int b = sector * T + t;
                if (cnt[b] == 0 || (used >> b & 1u)) {   // preferred residue dry: the fullest residue not yet used in this group
                    int best = -1;
                    for (int c = 0; c < 16; ++c)
                        if (!(used >> c & 1u) && cnt[c] > 0 && (best < 0 || cnt[c] > cnt[best])) best = c;
                    if (best < 0) {   // fewer distinct residues left than lanes: any entry
                        for (int c = 0; c < 16; ++c)
                            if (cnt[c] > 0 && (best < 0 || cnt[c] > cnt[best])) best = c;
                    }
                    b = best;
                }
                used |= 1u << b;
                put(g * T + t, byres[head[b]]);
                head[b]++;
                cnt[b]--;
                placed++;
            }
        }
    }
}

static int ilog2(int64_t v) {
    int l = 0;
    while ((1LL << (l + 1)) <= v) ++l;
    return l;
}

// Nonzeros a chunk should hold.  kSellTargetNnz on large panels (the staging of x amortises best); a panel with few
// chunks per resident block (a 1/8 slab of the 20M-dof system has ~1.4 chunks of B per block) gets smaller chunks so
// that the persistent grid's last round is not mostly idle.  NSB_SELL_TARGET_NNZ overrides (tuning sweeps).
static double sell_target_nnz(int64_t nnz, int num_sms) {
    const char* env = getenv("NSB_SELL_TARGET_NNZ");
    if (env && atof(env) >= 256.0) return atof(env);
    const char* per = getenv("NSB_SELL_CHUNKS_PER_BLOCK");
    const double chunks_per_block = per ? atof(per) : 0.0;   // 0: fixed target
    if (chunks_per_block <= 0.0) return kSellTargetNnz;
    double t = (double)nnz / ((double)num_sms * NSB_SELL_MINBLOCKS * chunks_per_block);
    if (t > kSellTargetNnz) t = kSellTargetNnz;
    if (t < 3000.0) t = 3000.0;
    return t;
}

static int pick_lanes_per_row(double mean) {
    const char* env = getenv("NSB_SELL_T");
    if (env) {
        int v = atoi(env);
        if (v == 1 || v == 2 || v == 4 || v == 8 || v == 16 || v == 32) return v;
    }
    // measured on the 20M-dof P2/P1 blocks (tools/sweep_sell.sh): A (28/row) best at 1, B (189/row) at 8
    if (mean <= 32.0) return 1;
    if (mean <= 64.0) return 2;
    if (mean <= 128.0) return 4;
    if (mean <= 256.0) return 8;
    if (mean <= 512.0) return 16;
    return 32;
}

__global__ void k_sell_identity_perm(unsigned short* __restrict__ perm, int64_t nvr, int mask) {
    for (int64_t i = (int64_t)blockIdx.x * blockDim.x + threadIdx.x; i < nvr; i += (int64_t)gridDim.x * blockDim.x)
        perm[i] = (unsigned short)(i & mask);
}

// A row panel without a single nonzero (the condensed-out rows of a projection operator, the V2 rows of bproj): the SELL kernel
// still has to run its row epilogue over these rows (y = 0, or xin + 0 for the fused adds), so the panel gets the smallest legal
// shape with empty slices -- per chunk the kernel reads 2 B per row of perm plus the slice offsets and stages nothing.
static int build_empty_panel(Context& c, SellData* ms, int64_t rows) {
    const int log2R = 8, T = 1;
    const int64_t R = 1LL << log2R;
    const int nchunks = (int)((rows + R - 1) / R);
    const int64_t nvr = (int64_t)nchunks * R, nslices = nvr / (32 / T);
    NSB_CHECK(ms->sell_perm.alloc(nvr));
    NSB_CHECK(ms->sell_off.alloc(nslices + 1));
    NSB_CHECK(ms->sell_uptr.alloc(nchunks + 1));
    NSB_CHECK(ms->sell_ucol.alloc(1));
    NSB_CHECK(ms->sell_vals.alloc(64));
    NSB_CHECK(ms->sell_lcol.alloc(64));
    k_sell_identity_perm<<<vec_grid(nvr, c.num_sms), kBlock, 0, c.stream>>>(ms->sell_perm.p, nvr, (int)R - 1);
    NSB_CUDA(cudaMemsetAsync(ms->sell_off.p, 0, sizeof(unsigned int) * (size_t)(nslices + 1), c.stream));
    NSB_CUDA(cudaMemsetAsync(ms->sell_uptr.p, 0, sizeof(int) * (size_t)(nchunks + 1), c.stream));
    NSB_CUDA(cudaMemsetAsync(ms->sell_ucol.p, 0, sizeof(int), c.stream));
    NSB_CUDA(cudaMemsetAsync(ms->sell_vals.p, 0, sizeof(double) * 64, c.stream));
    NSB_CUDA(cudaMemsetAsync(ms->sell_lcol.p, 0, sizeof(unsigned short) * 64, c.stream));
    NSB_CUDA(cudaStreamSynchronize(c.stream));
    ms->sell_R = (int)R;
    ms->sell_log2R = log2R;
    ms->sell_T = T;
    ms->sell_nchunks = nchunks;
    ms->sell_max_uniq = 0;
    ms->sell_cap = 32;
    ms->sell_total_uniq = 0;
    ms->sell_stored = 0;
    ms->sell_nruns = 0;
    return NSB_OK;
}

// Try to build the SELL-LX storage of the row panel [row0, row0 + rows) of m into S.  Leaves S.sell_R == 0 when the
// panel does not qualify.
static int build_sell_panel(Context& c, nsb_mat_s* m, int64_t row0, int64_t rows, SellData& S_) {
    SellData* ms = &S_;
    ms->sell_R = 0;
    ms->row0 = row0;
    ms->nrows = rows;
    const bool debug = getenv("NSB_SELL_DEBUG") != nullptr;
    if (rows < 1) return NSB_OK;
    int pends[2] = {0, 0};
    NSB_CUDA(cudaMemcpy(&pends[0], m->rowptr.p + row0, sizeof(int), cudaMemcpyDeviceToHost));
    NSB_CUDA(cudaMemcpy(&pends[1], m->rowptr.p + row0 + rows, sizeof(int), cudaMemcpyDeviceToHost));
    const int pbase = pends[0];
    const int64_t nnz = (int64_t)pends[1] - pends[0];
    if (nnz < 1) return build_empty_panel(c, ms, rows);
    const double mean = (double)nnz / (double)rows;
    const int T = pick_lanes_per_row(mean);
    const int rps = 32 / T;
    const int minlog = ilog2((int64_t)rps * kSellWarps);   // at least one slice per warp
    const double target_nnz = sell_target_nnz(nnz, c.num_sms);
    double want = target_nnz / (mean > 1.0 ? mean : 1.0);
    int log2R = ilog2((int64_t)(want * 1.41421356));        // nearest power of two
    const char* envR = getenv("NSB_SELL_LOG2R");
    if (envR) log2R = atoi(envR);
    if (log2R > 11) log2R = 11;
    if (log2R < minlog) log2R = minlog;
    CsrView A{rows, m->rowptr.p + row0, m->colidx.p, m->values.p};   // the panel's rows; positions stay absolute
    const int vg = vec_grid(nnz, c.num_sms);
    const int64_t stage_cap = getenv("NSB_SELL_STAGE") ? atoll(getenv("NSB_SELL_STAGE")) : kSellStage;
    const int64_t stage_max = stage_cap > kSellStageMax ? stage_cap : kSellStageMax;

    // ---- phase A: chunk-local column numbering (largest R whose chunks fit the staging buffer)
    DevBuf<unsigned short> lcol_csr;
    int nchunks = 0, max_uniq = 0, total_uniq = 0;
    bool found = false;
    {
        DevBuf<long long> k0, k1;
        DevBuf<int> p0, p1, head, gidx, dmax;
        DevBuf<char> tmp;
        NSB_CHECK(k0.alloc(nnz));
        NSB_CHECK(k1.alloc(nnz));
        NSB_CHECK(p0.alloc(nnz));
        NSB_CHECK(p1.alloc(nnz));
        NSB_CHECK(head.alloc(nnz));
        NSB_CHECK(gidx.alloc(nnz));
        NSB_CHECK(dmax.alloc(1));
        size_t sort_bytes = 0, scan_bytes = 0;
        cub::DeviceRadixSort::SortPairs(nullptr, sort_bytes, k0.p, k1.p, p0.p, p1.p, (int)nnz, 0, 64, c.stream);
        cub::DeviceScan::InclusiveSum(nullptr, scan_bytes, head.p, gidx.p, (int)nnz, c.stream);
        NSB_CHECK(tmp.alloc((int64_t)(sort_bytes > scan_bytes ? sort_bytes : scan_bytes) + 16));
        for (; log2R >= minlog && !found; --log2R) {
            const int64_t R = 1LL << log2R;
            nchunks = (int)((rows + R - 1) / R);
            const int end_bit = 32 + ilog2(nchunks) + 1;
            const int64_t kg = (rows + 7) / 8 < (int64_t)c.num_sms * 8 ? (rows + 7) / 8 : (int64_t)c.num_sms * 8;
            k_sell_keys<<<(int)kg, kBlock, 0, c.stream>>>(A, pbase, log2R, k0.p, p0.p);
            NSB_CUDA(cub::DeviceRadixSort::SortPairs(tmp.p, sort_bytes, k0.p, k1.p, p0.p, p1.p, (int)nnz, 0,
                                                     end_bit < 64 ? end_bit : 64, c.stream));
            k_sell_heads<<<vg, kBlock, 0, c.stream>>>(k1.p, head.p, nnz);
            NSB_CUDA(cub::DeviceScan::InclusiveSum(tmp.p, scan_bytes, head.p, gidx.p, (int)nnz, c.stream));
            DevBuf<int> uptr;
            NSB_CHECK(uptr.alloc(nchunks + 1));
            k_sell_uptr<<<(nchunks + 1 + kBlock - 1) / kBlock, kBlock, 0, c.stream>>>(k1.p, gidx.p, nnz, uptr.p, nchunks);
            NSB_CUDA(cudaMemsetAsync(dmax.p, 0, sizeof(int), c.stream));
            k_sell_maxuniq<<<(nchunks + kBlock - 1) / kBlock, kBlock, 0, c.stream>>>(uptr.p, nchunks, dmax.p);
            int h[2] = {0, 0};
            NSB_CUDA(cudaMemcpyAsync(&h[0], dmax.p, sizeof(int), cudaMemcpyDeviceToHost, c.stream));
            NSB_CUDA(cudaMemcpyAsync(&h[1], uptr.p + nchunks, sizeof(int), cudaMemcpyDeviceToHost, c.stream));
            NSB_CUDA(cudaStreamSynchronize(c.stream));
            if (debug)
                fprintf(stderr, "[sell] rows=%lld nnz=%lld T=%d log2R=%d nchunks=%d max_distinct=%d total=%d\n",
                        (long long)rows, (long long)nnz, T, log2R, nchunks, h[0], h[1]);
            // staging buffers (x + row sums) over budget: halve the chunk.  Chunks with fewer than two slices per
            // warp lose more to imbalance than to the larger footprint, so they may use the relaxed budget.
            const bool small = log2R <= ilog2((int64_t)rps * kSellWarps * 2);
            if (h[0] + R > (small ? stage_max : stage_cap)) continue;
            NSB_CHECK(lcol_csr.alloc(nnz));
            NSB_CHECK(ms->sell_ucol.alloc(h[1] > 0 ? h[1] : 1));
            k_sell_localcols<<<vg, kBlock, 0, c.stream>>>(k1.p, p1.p, head.p, gidx.p, uptr.p, lcol_csr.p, ms->sell_ucol.p, nnz);
            NSB_CUDA(cudaStreamSynchronize(c.stream));
            ms->sell_uptr.release();
            ms->sell_uptr.p = uptr.p;   // adopt
            ms->sell_uptr.n = uptr.n;
            uptr.p = nullptr;
            max_uniq = h[0];
            total_uniq = h[1];
            found = true;
            break;
        }
    }
    if (!found) {
        ms->sell_ucol.release();
        return NSB_OK;
    }
    // ---- runs of consecutive distinct columns
    ms->sell_nruns = 0;
    if (total_uniq > 0 && !(getenv("NSB_SELL_RUNS") && atoi(getenv("NSB_SELL_RUNS")) == 0)) {
        DevBuf<int> flag, ridx;
        DevBuf<char> stmp;
        NSB_CHECK(flag.alloc(total_uniq));
        NSB_CHECK(ridx.alloc(total_uniq));
        size_t sb = 0;
        cub::DeviceScan::InclusiveSum(nullptr, sb, flag.p, ridx.p, total_uniq, c.stream);
        NSB_CHECK(stmp.alloc((int64_t)sb + 16));
        const int ug = vec_grid(total_uniq, c.num_sms);
        NSB_CUDA(cudaMemsetAsync(flag.p, 0, sizeof(int) * (size_t)total_uniq, c.stream));
        k_sell_chunkflags<<<(nchunks + kBlock - 1) / kBlock, kBlock, 0, c.stream>>>(ms->sell_uptr.p, nchunks, total_uniq, flag.p);
        k_sell_runheads<<<ug, kBlock, 0, c.stream>>>(ms->sell_ucol.p, total_uniq, flag.p);
        NSB_CUDA(cub::DeviceScan::InclusiveSum(stmp.p, sb, flag.p, ridx.p, total_uniq, c.stream));
        int nruns = 0;
        NSB_CUDA(cudaMemcpyAsync(&nruns, ridx.p + (total_uniq - 1), sizeof(int), cudaMemcpyDeviceToHost, c.stream));
        NSB_CUDA(cudaStreamSynchronize(c.stream));
        if (debug) fprintf(stderr, "[sell] %d runs for %d distinct columns (mean run %.1f)\n", nruns, total_uniq, (double)total_uniq / nruns);
        if ((int64_t)nruns * 8 <= (int64_t)total_uniq) {   // mean run >= 8 columns: descriptors beat the list
            NSB_CHECK(ms->sell_rstart.alloc(nruns));
            NSB_CHECK(ms->sell_rpos.alloc(nruns + 1));
            NSB_CHECK(ms->sell_rptr.alloc(nchunks + 1));
            k_sell_runfill<<<ug, kBlock, 0, c.stream>>>(ms->sell_ucol.p, flag.p, ridx.p, total_uniq, ms->sell_rstart.p, ms->sell_rpos.p);
            k_sell_runptr<<<(nchunks + 1 + kBlock - 1) / kBlock, kBlock, 0, c.stream>>>(ms->sell_uptr.p, ridx.p, nchunks, total_uniq,
                                                                                         nruns, ms->sell_rptr.p, ms->sell_rpos.p);
            NSB_CUDA(cudaStreamSynchronize(c.stream));
            ms->sell_nruns = nruns;
        }
    }
    const int64_t R = 1LL << log2R;
    const int64_t nvr = (int64_t)nchunks * R;        // virtual rows (last chunk padded)
    const int64_t nslices = nvr / rps;

    // ---- phase B: sort the rows of each chunk by length, lay out the slices
    DevBuf<unsigned long long> rk0, rk1;
    DevBuf<unsigned int> npb;
    DevBuf<char> tmp;
    NSB_CHECK(rk0.alloc(nvr));
    NSB_CHECK(rk1.alloc(nvr));
    NSB_CHECK(npb.alloc(nslices + 1));
    NSB_CHECK(ms->sell_perm.alloc(nvr));
    NSB_CHECK(ms->sell_off.alloc(nslices + 1));
    size_t sort_bytes = 0, scan_bytes = 0;
    cub::DeviceRadixSort::SortKeys(nullptr, sort_bytes, rk0.p, rk1.p, (int)nvr, 0, 64, c.stream);
    cub::DeviceScan::ExclusiveSum(nullptr, scan_bytes, npb.p, ms->sell_off.p, (int)(nslices + 1), c.stream);
    NSB_CHECK(tmp.alloc((int64_t)(sort_bytes > scan_bytes ? sort_bytes : scan_bytes) + 16));
    const int rg = vec_grid(nvr, c.num_sms);
    k_sell_rowkeys<<<rg, kBlock, 0, c.stream>>>(A, log2R, nvr, rk0.p);
    const int end_bit = 40 + ilog2(nchunks) + 1;
    NSB_CUDA(cub::DeviceRadixSort::SortKeys(tmp.p, sort_bytes, rk0.p, rk1.p, (int)nvr, 0, end_bit < 64 ? end_bit : 64, c.stream));
    NSB_CUDA(cudaMemsetAsync(npb.p, 0, sizeof(unsigned int) * (size_t)(nslices + 1), c.stream));
    k_sell_slices<<<rg, kBlock, 0, c.stream>>>(rk1.p, nvr, rps, T, ms->sell_perm.p, npb.p);
    NSB_CUDA(cub::DeviceScan::ExclusiveSum(tmp.p, scan_bytes, npb.p, ms->sell_off.p, (int)(nslices + 1), c.stream));
    unsigned int total_pb = 0;
    NSB_CUDA(cudaMemcpyAsync(&total_pb, ms->sell_off.p + nslices, sizeof(unsigned int), cudaMemcpyDeviceToHost, c.stream));
    NSB_CUDA(cudaStreamSynchronize(c.stream));
    const int64_t stored = (int64_t)total_pb * 64;
    if (stored > 6 * nnz + 4096 || total_pb > 0x7FFFFFFFu / 2) {   // pathological padding: keep CSR-vector
        if (debug) fprintf(stderr, "[sell] rejected: stored=%lld for nnz=%lld\n", (long long)stored, (long long)nnz);
        ms->sell_perm.release();
        ms->sell_off.release();
        ms->sell_uptr.release();
        ms->sell_ucol.release();
        return NSB_OK;
    }
    NSB_CHECK(ms->sell_vals.alloc(stored > 0 ? stored : 64));
    NSB_CHECK(ms->sell_lcol.alloc(stored > 0 ? stored : 64));
    NSB_CUDA(cudaMemsetAsync(ms->sell_vals.p, 0, sizeof(double) * (size_t)(stored > 0 ? stored : 64), c.stream));
    NSB_CUDA(cudaMemsetAsync(ms->sell_lcol.p, 0, sizeof(unsigned short) * (size_t)(stored > 0 ? stored : 64), c.stream));
    const char* envb = getenv("NSB_SELL_BANKED");   // 1: bank-aware entry order inside the rows of T >= 4 panels (see k_sell_fill_banked)
    const bool banked = T >= 4 && T <= 16 && envb && atoi(envb) != 0;
    if (banked) {
        const int64_t bg = (nvr + 127) / 128 < (int64_t)c.num_sms * 16 ? (nvr + 127) / 128 : (int64_t)c.num_sms * 16;
        k_sell_fill_banked<<<(int)bg, 128, 0, c.stream>>>(A, pbase, lcol_csr.p, log2R, nvr, rps, T, ms->sell_perm.p, ms->sell_off.p,
                                                          ms->sell_vals.p, ms->sell_lcol.p);
    } else {
        const int64_t fg = (nvr + 7) / 8 < (int64_t)c.num_sms * 8 ? (nvr + 7) / 8 : (int64_t)c.num_sms * 8;
        k_sell_fill<<<(int)fg, kBlock, 0, c.stream>>>(A, pbase, lcol_csr.p, log2R, nvr, rps, T, ms->sell_perm.p, ms->sell_off.p,
                                                     ms->sell_vals.p, ms->sell_lcol.p);
    }
    NSB_CUDA(cudaGetLastError());
    NSB_CUDA(cudaStreamSynchronize(c.stream));
    ms->sell_R = (int)R;
    ms->sell_log2R = log2R;
    ms->sell_T = T;
    ms->sell_nchunks = nchunks;
    ms->sell_max_uniq = max_uniq;
    ms->sell_cap = (max_uniq + 31) & ~31;
    if (ms->sell_cap < 32) ms->sell_cap = 32;
    ms->sell_total_uniq = total_uniq;
    ms->sell_stored = stored;
    if (debug)
        fprintf(stderr, "[sell] built: R=%lld T=%d chunks=%d stored=%lld (%.2f%% padding) smem=%zu B\n", (long long)R, T, nchunks,
                (long long)stored, 100.0 * (double)(stored - nnz) / (double)nnz, sell_smem_bytes(ms));
    return NSB_OK;
}

// ---- interior-first chunk order of a row-partitioned operator ---------------------------------------------
// A chunk needs the halo exchange iff its largest distinct column is a halo column (>= n_owned; the lists are ascending).
__global__ void k_sell_halo_flags(const int* __restrict__ uptr, const int* __restrict__ ucol, int nchunks, int64_t n_owned,
                                  int* __restrict__ flag) {
    for (int c = blockIdx.x * blockDim.x + threadIdx.x; c < nchunks; c += gridDim.x * blockDim.x) {
        const int u0 = uptr[c], u1 = uptr[c + 1];
        flag[c] = (u1 > u0 && (int64_t)ucol[u1 - 1] >= n_owned) ? 1 : 0;
    }
}

int build_halo_order(Context& c, nsb_mat_s* m) {
    for (auto& P : m->panels) {
        P->halo_n_interior = -1;
        P->halo_order.release();
    }
    if (!m->halo || m->panels.empty()) return NSB_OK;
    for (auto& P : m->panels) {
        const int nc = P->sell_nchunks;
        if (nc < 1) continue;
        DevBuf<int> flag;
        NSB_CHECK(flag.alloc(nc));
        k_sell_halo_flags<<<(nc + kBlock - 1) / kBlock, kBlock, 0, c.stream>>>(P->sell_uptr.p, P->sell_ucol.p, nc, m->halo->n_owned, flag.p);
        NSB_CUDA(cudaGetLastError());
        std::vector<int> hf(nc), order(nc);
        NSB_CUDA(cudaMemcpyAsync(hf.data(), flag.p, sizeof(int) * (size_t)nc, cudaMemcpyDeviceToHost, c.stream));
        NSB_CUDA(cudaStreamSynchronize(c.stream));
        int k = 0;
        for (int i = 0; i < nc; ++i)
            if (!hf[i]) order[k++] = i;
        const int n_int = k;
        for (int i = 0; i < nc; ++i)
            if (hf[i]) order[k++] = i;
        NSB_CHECK(P->halo_order.alloc(nc));
        NSB_CUDA(cudaMemcpy(P->halo_order.p, order.data(), sizeof(int) * (size_t)nc, cudaMemcpyHostToDevice));
        P->halo_n_interior = n_int;
        if (getenv("NSB_SELL_DEBUG"))
            fprintf(stderr, "[sell] halo order: %d of %d chunks touch halo columns\n", nc - n_int, nc);
    }
    return NSB_OK;
}

// Build the SELL-LX companion of m, one panel per entry of m->panel_splits (default: the whole matrix).  Leaves
// m->panels empty when the operator is small or any panel does not qualify (the CSR-vector kernel then runs).
int build_sell(Context& c, nsb_mat_s* m) {
    NvtxRange nvtx_call("build_sell (SELL-LX companion)");
    m->panels.clear();
    const char* env = getenv("NSB_SELL");
    if (env && atoi(env) == 0) return NSB_OK;
    const int64_t min_nnz = getenv("NSB_SELL_MIN_NNZ") ? atoll(getenv("NSB_SELL_MIN_NNZ")) : 200000;
    if (m->nnz < min_nnz || m->rows < 1 || m->nnz < 1) return NSB_OK;
    std::vector<int64_t> sp = m->panel_splits;
    if (sp.size() < 2 || sp.front() != 0 || sp.back() != m->rows) sp = {0, m->rows};
    for (size_t k = 0; k + 1 < sp.size(); ++k) {
        if (sp[k + 1] <= sp[k]) continue;
        std::unique_ptr<SellData> S(new SellData());
        int r = build_sell_panel(c, m, sp[k], sp[k + 1] - sp[k], *S);
        if (r != NSB_OK) {
            m->panels.clear();
            return r;
        }
        if (S->sell_R == 0) {   // a panel that does not fit sends the whole operator to the CSR-vector kernel
            m->panels.clear();
            return NSB_OK;
        }
        m->panels.push_back(std::move(S));
    }
    return m->halo ? build_halo_order(c, m) : NSB_OK;
}

}  // namespace nsb
