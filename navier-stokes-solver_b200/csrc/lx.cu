// lx.cu -- builds the chunk-local column index ("LX") companion of a CSR operator.
//
// ncu on the plain CSR-vector kernel (profiles/r1_spmv_csrvec_A_2Mdofs_ncu.csv) shows DRAM traffic equal to
// the algorithmic bytes but the L1 data pipe at 72 %: every warp-wide gather of x touches 5+ cache
// lines, so the SM's L1, not HBM, bounds the kernel at ~60 % of the copy roofline.  The fix is the
// "shared-memory staging of x" of the north star, made general:
//   * rows are cut into chunks of R consecutive rows;
//   * per chunk the sorted set of distinct columns is stored once (ucol, 4 B each);
//   * every nonzero stores a 16-bit index into that set (lcol) instead of a 32-bit global column.
// The SpMV kernel (spmv.cuh: spmv_lx_kernel) stages x[ucol] of a chunk into shared memory with
// coalesced loads and then gathers from shared memory, so gathers cost bank cycles instead of cache
// lines, and the index stream shrinks from 4 to 2 bytes per nonzero.  An operator whose chunks need
// more than kLxCap distinct columns even at the smallest R keeps the CSR-vector kernel.
#include <cub/cub.cuh>
#include <stdlib.h>

#include "spmv.cuh"

namespace nsb {

__global__ void k_lx_keys(CsrView A, int log2R, long long* __restrict__ keys, int* __restrict__ pos) {
    const int lane = threadIdx.x & 31;
    const int64_t warp = ((int64_t)blockIdx.x * blockDim.x + threadIdx.x) >> 5;
    const int64_t nwarps = ((int64_t)gridDim.x * blockDim.x) >> 5;
    for (int64_t r = warp; r < A.rows; r += nwarps) {
        const int p0 = A.rowptr[r], p1 = A.rowptr[r + 1];
        const long long hi = (long long)(r >> log2R) << 32;
        for (int p = p0 + lane; p < p1; p += 32) {
            keys[p] = hi | (long long)A.colidx[p];
            pos[p] = p;
        }
    }
}

__global__ void k_lx_heads(const long long* __restrict__ keys, int* __restrict__ head, int64_t nnz) {
    for (int64_t i = (int64_t)blockIdx.x * blockDim.x + threadIdx.x; i < nnz; i += (int64_t)gridDim.x * blockDim.x)
        head[i] = (i == 0 || keys[i] != keys[i - 1]) ? 1 : 0;
}

// uptr[c] = number of distinct (chunk, col) pairs with chunk < c  = gidx at the first sorted entry of chunk >= c
__global__ void k_lx_uptr(const long long* __restrict__ keys, const int* __restrict__ gidx_incl, int64_t nnz,
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

__global__ void k_lx_scatter(const long long* __restrict__ keys, const int* __restrict__ pos,
                             const int* __restrict__ head, const int* __restrict__ gidx_incl,
                             const int* __restrict__ uptr, unsigned short* __restrict__ lcol, int* __restrict__ ucol,
                             int64_t nnz) {
    for (int64_t i = (int64_t)blockIdx.x * blockDim.x + threadIdx.x; i < nnz; i += (int64_t)gridDim.x * blockDim.x) {
        const long long k = keys[i];
        const int chunk = (int)(k >> 32);
        const int g = gidx_incl[i] - 1;           // global distinct index of this entry's column
        lcol[pos[i]] = (unsigned short)(g - uptr[chunk]);
        if (head[i]) ucol[g] = (int)(k & 0xffffffffLL);
    }
}

__global__ void k_lx_maxuniq(const int* __restrict__ uptr, int nchunks, int* __restrict__ out) {
    int m = 0;
    for (int c = blockIdx.x * blockDim.x + threadIdx.x; c < nchunks; c += gridDim.x * blockDim.x)
        m = max(m, uptr[c + 1] - uptr[c]);
    atomicMax(out, m);
}

static int ilog2(int64_t v) {
    int l = 0;
    while ((1LL << (l + 1)) <= v) ++l;
    return l;
}

// Try to build the LX companion of m.  Leaves m->lx_R == 0 when the operator does not qualify.
int build_lx(Context& c, nsb_mat_s* m) {
    m->lx_R = 0;
    const char* env = getenv("NSB_LX");
    if (env && atoi(env) == 0) return NSB_OK;
    const int64_t nnz = m->nnz, rows = m->rows;
    const int64_t min_nnz = getenv("NSB_LX_MIN_NNZ") ? atoll(getenv("NSB_LX_MIN_NNZ")) : 200000;
    if (getenv("NSB_LX_DEBUG")) fprintf(stderr, "[lx] consider rows=%lld nnz=%lld min_nnz=%lld\n", (long long)rows, (long long)nnz, (long long)min_nnz);
    if (nnz < min_nnz || rows < 1 || nnz < 1) return NSB_OK;
    const double mean = (double)nnz / (double)rows;
    const int tpr = m->threads_per_row;
    const int gpb = kLxBlock / tpr;                 // rows a block handles per pass
    int log2R = ilog2((int64_t)(kLxTargetNnz / (mean > 1.0 ? mean : 1.0)));
    const char* envR = getenv("NSB_LX_LOG2R");
    if (envR) log2R = atoi(envR);
    const int minlog = ilog2(gpb);
    if (log2R > 12) log2R = 12;
    if (log2R < minlog) log2R = minlog;

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
    const int vg = vec_grid(nnz, c.num_sms);
    CsrView A = view_of(m);

    for (; log2R >= minlog; --log2R) {
        const int64_t R = 1LL << log2R;
        const int nchunks = (int)((rows + R - 1) / R);
        const int end_bit = 32 + ilog2(nchunks) + 1;
        k_lx_keys<<<(int)((rows + 7) / 8 < c.num_sms * 8 ? (rows + 7) / 8 : c.num_sms * 8), kBlock, 0, c.stream>>>(A, log2R, k0.p, p0.p);
        NSB_CUDA(cub::DeviceRadixSort::SortPairs(tmp.p, sort_bytes, k0.p, k1.p, p0.p, p1.p, (int)nnz, 0,
                                                 end_bit < 64 ? end_bit : 64, c.stream));
        k_lx_heads<<<vg, kBlock, 0, c.stream>>>(k1.p, head.p, nnz);
        NSB_CUDA(cub::DeviceScan::InclusiveSum(tmp.p, scan_bytes, head.p, gidx.p, (int)nnz, c.stream));
        DevBuf<int> uptr;
        NSB_CHECK(uptr.alloc(nchunks + 1));
        k_lx_uptr<<<(nchunks + 1 + kBlock - 1) / kBlock, kBlock, 0, c.stream>>>(k1.p, gidx.p, nnz, uptr.p, nchunks);
        NSB_CUDA(cudaMemsetAsync(dmax.p, 0, sizeof(int), c.stream));
        k_lx_maxuniq<<<(nchunks + kBlock - 1) / kBlock, kBlock, 0, c.stream>>>(uptr.p, nchunks, dmax.p);
        int h[2] = {0, 0};
        NSB_CUDA(cudaMemcpyAsync(&h[0], dmax.p, sizeof(int), cudaMemcpyDeviceToHost, c.stream));
        NSB_CUDA(cudaMemcpyAsync(&h[1], uptr.p + nchunks, sizeof(int), cudaMemcpyDeviceToHost, c.stream));
        NSB_CUDA(cudaStreamSynchronize(c.stream));
        if (getenv("NSB_LX_DEBUG"))
            fprintf(stderr, "[lx] rows=%lld nnz=%lld tpr=%d log2R=%d nchunks=%d max_distinct=%d total=%d\n", (long long)rows,
                    (long long)nnz, tpr, log2R, nchunks, h[0], h[1]);
        if (h[0] > kLxCap) continue;  // too many distinct columns for the staging buffer: halve the chunk
        const int64_t padded = ((nnz + 1) & ~1LL) + 8;
        NSB_CHECK(m->lx_lcol.alloc(padded));
        NSB_CHECK(m->lx_ucol.alloc(h[1] > 0 ? h[1] : 1));
        NSB_CUDA(cudaMemsetAsync(m->lx_lcol.p, 0, sizeof(unsigned short) * (size_t)padded, c.stream));
        k_lx_scatter<<<vg, kBlock, 0, c.stream>>>(k1.p, p1.p, head.p, gidx.p, uptr.p, m->lx_lcol.p, m->lx_ucol.p, nnz);
        NSB_CUDA(cudaStreamSynchronize(c.stream));
        m->lx_uptr.p = uptr.p;   // adopt
        m->lx_uptr.n = uptr.n;
        uptr.p = nullptr;
        m->lx_R = (int)R;
        m->lx_nchunks = nchunks;
        m->lx_max_uniq = h[0];
        m->lx_total_uniq = h[1];
        return NSB_OK;
    }
    return NSB_OK;
}

}  // namespace nsb
