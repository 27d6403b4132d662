// vecops.cu -- unfused vector primitives for the set-up phases of the solvers (initial residuals,
// transformed right-hand sides).  The iteration bodies use the fused kernels in minres.cu /
// bpcg.cu / cg.cu instead.  Replaces NGSolve vector expressions `x.data = a*y + b*z`, InnerProduct
// (e.g. minres.py:64-74, solvers/bramblepasciak_new.py:124-189).
#include "common.cuh"

namespace nsb {

__global__ void __launch_bounds__(kBlock) k_set(double* __restrict__ y, double a, int64_t n) {
    for (int64_t i = (int64_t)blockIdx.x * blockDim.x + threadIdx.x; i < n; i += (int64_t)gridDim.x * blockDim.x) y[i] = a;
}

__global__ void __launch_bounds__(kBlock) k_lincomb(double* y, double a, const double* x1, double b, const double* x2,
                                                    double cc, const double* x3, int64_t n) {
    for (int64_t i = (int64_t)blockIdx.x * blockDim.x + threadIdx.x; i < n; i += (int64_t)gridDim.x * blockDim.x) {
        double v = a * x1[i];
        if (x2) v += b * x2[i];
        if (x3) v += cc * x3[i];
        y[i] = v;
    }
}

__global__ void __launch_bounds__(kBlock) k_diagmul(double* y, double s, const double* d, const double* x, int64_t n) {
    for (int64_t i = (int64_t)blockIdx.x * blockDim.x + threadIdx.x; i < n; i += (int64_t)gridDim.x * blockDim.x)
        y[i] = d ? s * (d[i] * x[i]) : s * x[i];
}

__global__ void __launch_bounds__(kBlock) k_dot(const double* __restrict__ x, const double* __restrict__ y, int64_t n,
                                                double* out, double* partials, unsigned int* ticket) {
    double acc[1] = {0.0};
    for (int64_t i = (int64_t)blockIdx.x * blockDim.x + threadIdx.x; i < n; i += (int64_t)gridDim.x * blockDim.x)
        acc[0] = fma(x[i], y[i], acc[0]);
    if (grid_sum<1>(acc, partials, ticket)) {
        if (threadIdx.x == 0) *out = acc[0];
    }
}

int vec_set(Context& c, double* y, double a, int64_t n) {
    if (n <= 0) return NSB_OK;
    k_set<<<vec_grid(n, c.num_sms), kBlock, 0, c.stream>>>(y, a, n);
    c.launches++;
    NSB_CUDA(cudaGetLastError());
    return NSB_OK;
}

int vec_lincomb(Context& c, double* y, double a, const double* x1, double b, const double* x2, double cc,
                const double* x3, int64_t n) {
    if (n <= 0) return NSB_OK;
    k_lincomb<<<vec_grid(n, c.num_sms), kBlock, 0, c.stream>>>(y, a, x1, b, x2, cc, x3, n);
    c.launches++;
    NSB_CUDA(cudaGetLastError());
    return NSB_OK;
}

int vec_diagmul(Context& c, double* y, double s, const double* d, const double* x, int64_t n) {
    if (n <= 0) return NSB_OK;
    k_diagmul<<<vec_grid(n, c.num_sms), kBlock, 0, c.stream>>>(y, s, d, x, n);
    c.launches++;
    NSB_CUDA(cudaGetLastError());
    return NSB_OK;
}

int vec_dot(Context& c, const double* x, const double* y, int64_t n, double* out_dev) {
    k_dot<<<vec_grid(n > 0 ? n : 1, c.num_sms), kBlock, 0, c.stream>>>(x, y, n, out_dev, c.partials, c.ticket);
    c.launches++;
    NSB_CUDA(cudaGetLastError());
    return allreduce_sum(c, out_dev, 1);
}

}  // namespace nsb

// ---- device-resident vector helpers of the time-stepper chains (templates/NavierStokesSIMPLE_iterative.py:427-443) ----
namespace nsb {

__global__ void __launch_bounds__(kBlock) k_axpby(double* __restrict__ y, double a, const double* __restrict__ x, double b, int64_t n) {
    for (int64_t i = (int64_t)blockIdx.x * blockDim.x + threadIdx.x; i < n; i += (int64_t)gridDim.x * blockDim.x)
        y[i] = b == 0.0 ? a * x[i] : a * x[i] + b * y[i];
}

// y[i] += a * x[idx[i]]   (PermutationMatrix / Embedding apply: vel.data -= mapV * tmp_projection, :442)
__global__ void __launch_bounds__(kBlock) k_gather_axpy(double* __restrict__ y, double a, const double* __restrict__ x,
                                                        const int* __restrict__ idx, int64_t n) {
    for (int64_t i = (int64_t)blockIdx.x * blockDim.x + threadIdx.x; i < n; i += (int64_t)gridDim.x * blockDim.x)
        y[i] += a * x[idx[i]];
}

}  // namespace nsb

using namespace nsb;

extern "C" int nsb_vec_axpby(nsb_ctx ctx, double* y, double a, const double* x, double b, int64_t n) {
    NSB_REQUIRE(ctx && y && x && n >= 0, "nsb_vec_axpby: bad argument");
    NSB_REQUIRE(is_device_ptr(y) && is_device_ptr(x), "nsb_vec_axpby: device-resident vectors only");
    Context& c = ctx->c;
    NSB_CUDA(cudaSetDevice(c.device));
    if (n == 0) return NSB_OK;
    k_axpby<<<vec_grid(n, c.num_sms), kBlock, 0, c.stream>>>(y, a, x, b, n);
    c.launches++;
    NSB_CUDA(cudaGetLastError());
    NSB_CUDA(cudaStreamSynchronize(c.stream));
    return NSB_OK;
}

extern "C" int nsb_vec_gather_axpy(nsb_ctx ctx, double* y, double a, const double* x, const int32_t* idx, int64_t n) {
    NSB_REQUIRE(ctx && y && x && idx && n >= 0, "nsb_vec_gather_axpy: bad argument");
    NSB_REQUIRE(is_device_ptr(y) && is_device_ptr(x) && is_device_ptr(idx), "nsb_vec_gather_axpy: device-resident arrays only");
    Context& c = ctx->c;
    NSB_CUDA(cudaSetDevice(c.device));
    if (n == 0) return NSB_OK;
    k_gather_axpy<<<vec_grid(n, c.num_sms), kBlock, 0, c.stream>>>(y, a, x, idx, n);
    c.launches++;
    NSB_CUDA(cudaGetLastError());
    NSB_CUDA(cudaStreamSynchronize(c.stream));
    return NSB_OK;
}
