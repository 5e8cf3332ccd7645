// K7 / K8 — batched scalar samplers.
//
//  * sq_fejer_sample: one outcome per element of the amplitude / phase
//    estimation law (amplitude_estimation, Utility.py:494-508;
//    phase_estimation, Utility.py:637-653) in O(1) expected time, see
//    fejer.cuh.  sq_fejer_sample_bruteforce is the O(M) inversion the
//    reference performs; it exists as the device-side truth for tests.
//  * sq_truncnorm_add: x + TruncNormal(0, 1; [-a, a]) per element
//    (make_gaussian_est, Utility.py:88-104; introduce_error(_array), :68-73,
//    all through scipy.stats.truncnorm.rvs), by inverse-CDF.
//  * sq_binomial_sample: test hook for the binomial sampler of binomial.cuh.
#include "binomial.cuh"
#include "common.cuh"
#include "sqb200.h"
#include "fejer.cuh"

namespace sq {

__global__ void fejer_kernel(const double* __restrict__ M, const double* __restrict__ x, long long B,
                             unsigned long long seed, unsigned long long offset, double* __restrict__ out) {
    long long i = (long long)blockIdx.x * blockDim.x + threadIdx.x;
    if (i >= B) return;
    Philox rng(seed, offset + (unsigned long long)i, 0xFE7Eu, 0u);
    out[i] = fejer_sample(M[i], x[i], rng);
}

__global__ void fejer_brute_kernel(const double* __restrict__ M, const double* __restrict__ x, long long B,
                                   unsigned long long seed, unsigned long long offset, double* __restrict__ out) {
    long long i = (long long)blockIdx.x * blockDim.x + threadIdx.x;
    if (i >= B) return;
    Philox rng(seed, offset + (unsigned long long)i, 0xB2F0u, 0u);
    double Mi = M[i], xi = x[i];
    double j0 = floor(xi), t = xi - j0;
    if (t == 0.0) { out[i] = j0 >= Mi ? j0 - Mi : j0; return; }
    double st = sinpi(t), s2 = st * st;
    double u = rng.uniform(), acc = 0.0, pick = Mi - 1.0;
    for (double j = 0.0; j < Mi; j += 1.0) {
        double s = Mi * sinpi((j - xi) / Mi);
        acc += s2 / (s * s);
        if (u < acc) { pick = j; break; }
    }
    out[i] = pick;
}

__global__ void truncnorm_add_kernel(const double* __restrict__ x, long long L, double a, unsigned long long seed,
                                     unsigned long long offset, double* __restrict__ out) {
    long long i = (long long)blockIdx.x * blockDim.x + threadIdx.x;
    if (i >= L) return;
    Philox rng(seed, offset + (unsigned long long)i, 0x7A0Cu, 0u);
    double u = rng.uniform();
    double lo = normcdf(-a), hi = normcdf(a);
    double z = normcdfinv(lo + u * (hi - lo));
    z = fmin(fmax(z, -a), a);
    out[i] = x[i] + z;
}

__global__ void binomial_kernel(const double* __restrict__ n, const double* __restrict__ pl, const double* __restrict__ pr,
                                long long B, unsigned long long seed, unsigned long long offset, double* __restrict__ out) {
    long long i = (long long)blockIdx.x * blockDim.x + threadIdx.x;
    if (i >= B) return;
    Philox rng(seed, offset + (unsigned long long)i, 0xB100u, 0u);
    out[i] = binomial_split(n[i], pl[i], pr[i], rng);
}

}  // namespace sq

using namespace sq;

extern "C" int sq_fejer_sample(const double* M, const double* x, long long B, unsigned long long seed,
                               unsigned long long offset, double* out_j, void* stream_) {
    cudaStream_t stream = (cudaStream_t)stream_;
    SQ_REQUIRE(M && x && out_j && B >= 0, "sq_fejer_sample: bad arguments");
    if (B == 0) return SQ_OK;
    fejer_kernel<<<(unsigned)((B + 127) / 128), 128, 0, stream>>>(M, x, B, seed, offset, out_j);
    SQ_LAUNCH_CHECK("fejer_kernel");
    return SQ_OK;
}

extern "C" int sq_fejer_sample_bruteforce(const double* M, const double* x, long long B, unsigned long long seed,
                                          unsigned long long offset, double* out_j, void* stream_) {
    cudaStream_t stream = (cudaStream_t)stream_;
    SQ_REQUIRE(M && x && out_j && B >= 0, "sq_fejer_sample_bruteforce: bad arguments");
    if (B == 0) return SQ_OK;
    fejer_brute_kernel<<<(unsigned)((B + 127) / 128), 128, 0, stream>>>(M, x, B, seed, offset, out_j);
    SQ_LAUNCH_CHECK("fejer_brute_kernel");
    return SQ_OK;
}

extern "C" int sq_truncnorm_add(const double* x, long long L, double halfwidth, unsigned long long seed,
                                unsigned long long offset, double* out, void* stream_) {
    cudaStream_t stream = (cudaStream_t)stream_;
    SQ_REQUIRE(x && out && L >= 0, "sq_truncnorm_add: bad arguments");
    SQ_REQUIRE(halfwidth > 0.0, "sq_truncnorm_add: halfwidth must be > 0");
    if (L == 0) return SQ_OK;
    truncnorm_add_kernel<<<(unsigned)((L + 255) / 256), 256, 0, stream>>>(x, L, halfwidth, seed, offset, out);
    SQ_LAUNCH_CHECK("truncnorm_add_kernel");
    return SQ_OK;
}

extern "C" int sq_binomial_sample(const double* n, const double* pl, const double* pr, long long B,
                                  unsigned long long seed, unsigned long long offset, double* out, void* stream_) {
    cudaStream_t stream = (cudaStream_t)stream_;
    SQ_REQUIRE(n && pl && pr && out && B >= 0, "sq_binomial_sample: bad arguments");
    if (B == 0) return SQ_OK;
    binomial_kernel<<<(unsigned)((B + 127) / 128), 128, 0, stream>>>(n, pl, pr, B, seed, offset, out);
    SQ_LAUNCH_CHECK("binomial_kernel");
    return SQ_OK;
}
