// QLSSVC (sklearn/svm/_qSVM.py): kernel matrices on the fp64 tensor cores and per-sample error injection.
//
//  * sq_kernel_matrix: K = k(X, Y) for the four kernels get_kernel offers (_qSVM.py:375-389, over
//    metrics/pairwise.py): the n x m dot products come from the DMMA projection kernel (sq_project), the
//    kernel function is applied in place.
//  * sq_truncnorm_add_each: x_i + TruncNormal(0, 1; [-a_i, a_i]) with one half-width per element - the
//    error model of predict / get_P / get_h (_qSVM.py:199-214, 263-298: introduce_error(p, err / (2 beta_i))).
#include "common.cuh"
#include "sqb200.h"

namespace sq {

enum SvmKernel { SVM_LINEAR = 0, SVM_POLY = 1, SVM_RBF = 2, SVM_SIGMOID = 3 };

__global__ void svm_kernel_transform(double* __restrict__ K, long long n, int m, const double* __restrict__ xn,
                                     const double* __restrict__ yn, int kind, double gamma, double coef0,
                                     double degree, int same) {
    const long long total = n * (long long)m;
    for (long long idx = (long long)blockIdx.x * blockDim.x + threadIdx.x; idx < total;
         idx += (long long)gridDim.x * blockDim.x) {
        const long long i = idx / m;
        const int j = (int)(idx - i * m);
        const double g = K[idx];
        double v;
        if (kind == SVM_POLY) v = pow(gamma * g + coef0, degree);
        else if (kind == SVM_SIGMOID) v = tanh(gamma * g + coef0);
        else {
            // euclidean_distances(squared=True): |x|^2 + |y|^2 - 2 x.y clipped at 0, exact zeros on the diagonal of k(X, X)
            double d2 = fmax(xn[i] + yn[j] - 2.0 * g, 0.0);
            if (same && i == j) d2 = 0.0;
            v = exp(-gamma * d2);
        }
        K[idx] = v;
    }
}

__global__ void truncnorm_add_each_kernel(const double* __restrict__ x, long long L, const double* __restrict__ a,
                                          unsigned long long seed, unsigned long long offset,
                                          double* __restrict__ out) {
    long long i = (long long)blockIdx.x * blockDim.x + threadIdx.x;
    if (i >= L) return;
    const double ai = a[i];
    if (!(ai > 0.0)) { out[i] = x[i]; return; }
    Philox rng(seed, offset + (unsigned long long)i, 0x7A0Cu, 0u);
    double u = rng.uniform();
    double lo = normcdf(-ai), hi = normcdf(ai);
    double z = normcdfinv(lo + u * (hi - lo));
    z = fmin(fmax(z, -ai), ai);
    out[i] = x[i] + z;
}

}  // namespace sq

using namespace sq;

extern "C" size_t sq_kernel_matrix_workspace_bytes(long long n, int m) {
    return align_up((size_t)m * sizeof(double), 256) + align_up((size_t)n * sizeof(double), 256) +
           align_up((size_t)m * sizeof(double), 256);
}

extern "C" int sq_kernel_matrix(const double* X, long long n, const double* Y, int m, int d, int kind, double gamma,
                                double coef0, double degree, double* K, void* ws, size_t ws_bytes, void* stream_) {
    cudaStream_t stream = (cudaStream_t)stream_;
    SQ_REQUIRE(X && Y && K, "sq_kernel_matrix: null pointer");
    SQ_REQUIRE(n >= 1 && m >= 1 && d >= 1, "sq_kernel_matrix: bad shape n=%lld m=%d d=%d", n, m, d);
    SQ_REQUIRE(kind >= SVM_LINEAR && kind <= SVM_SIGMOID, "sq_kernel_matrix: kind must be 0 (linear), 1 (poly), 2 (rbf) or 3 (sigmoid)");
    const size_t need = sq_kernel_matrix_workspace_bytes(n, m);
    if (!ws || ws_bytes < need) { set_error("sq_kernel_matrix: workspace needs %zu bytes, got %zu", need, ws_bytes); return SQ_ERR_WORKSPACE; }
    char* w = (char*)ws;
    double* zero = (double*)w;
    double* xn = (double*)(w + align_up((size_t)m * sizeof(double), 256));
    double* yn = (double*)((char*)xn + align_up((size_t)n * sizeof(double), 256));
    SQ_CUDA(cudaMemsetAsync(zero, 0, (size_t)m * sizeof(double), stream));
    int rc = sq_project(X, n, d, Y, m, zero, K, stream_);
    if (rc) return rc;
    if (kind == SVM_LINEAR) return SQ_OK;
    if (kind == SVM_RBF) {
        rc = sq_row_sqnorms(X, n, d, xn, stream_);
        if (rc) return rc;
        rc = sq_row_sqnorms(Y, m, d, yn, stream_);
        if (rc) return rc;
    }
    const long long total = n * (long long)m;
    long long blocks = (total + 255) / 256;
    const int cap = device_sm_count() * 16;
    if (blocks > cap) blocks = cap;
    svm_kernel_transform<<<(unsigned)blocks, 256, 0, stream>>>(K, n, m, xn, yn, kind, gamma, coef0, degree, X == Y ? 1 : 0);
    SQ_LAUNCH_CHECK("svm_kernel_transform");
    return SQ_OK;
}

extern "C" int sq_truncnorm_add_each(const double* x, long long L, const double* halfwidth, unsigned long long seed,
                                     unsigned long long offset, double* out, void* stream_) {
    cudaStream_t stream = (cudaStream_t)stream_;
    SQ_REQUIRE(x && halfwidth && out && L >= 0, "sq_truncnorm_add_each: bad arguments");
    if (L == 0) return SQ_OK;
    truncnorm_add_each_kernel<<<(unsigned)((L + 255) / 256), 256, 0, stream>>>(x, L, halfwidth, seed, offset, out);
    SQ_LAUNCH_CHECK("truncnorm_add_each_kernel");
    return SQ_OK;
}
