// Stand-alone rate of the lean IPE estimator (csrc/ipe.cuh) at a chosen register cap / occupancy:
//   nvcc -gencode arch=compute_100a,code=sm_100a -O3 -std=c++17 -Isq-learn_b200/csrc -Iinclude -maxrregcount 64 -cudart shared -o ipe_mb tools/ipe_microbench.cu
//   ./ipe_mb <threads per CTA> <CTAs per SM>
// Measured on B200: 12.7 / 11.8 / 12.4 Gpairs/s at 128 / 80 / 64 registers (16 / 24 / 32 warps per SM): occupancy does not move it.
#include "ipe.cuh"
using namespace sq;
__global__ void k(const double* dot, const double* xn, const double* cn, double* out, int* flag, long long n, int kk,
                  double eps, unsigned long long seed) {
    for (long long i0 = (long long)blockIdx.x * blockDim.x; i0 < n; i0 += (long long)gridDim.x * blockDim.x) {
        long long i = i0 + threadIdx.x;
        bool ok = i < n;
        long long row = ok ? i / kk : 0; int col = ok ? (int)(i - row * kk) : 0;
        bool defer;
        double v = ipe_estimate_warp<true>(ok ? dot[i] : 0.0, xn[row], cn[col], eps, ok, seed, (unsigned long long)row, (unsigned)col, false, defer);
        if (ok) { out[i] = v; flag[i] = defer; }
    }
}

#include <cstdio>
#include <cstdlib>
#include <vector>
#include <cmath>
int main(int argc, char** argv) {
    const long long rows = 2000000; const int kk = 256; const long long n = rows * kk;
    int threads = argc > 1 ? atoi(argv[1]) : 256;
    int per_sm = argc > 2 ? atoi(argv[2]) : 2;
    std::vector<double> hx(rows), hc(kk);
    for (long long i = 0; i < rows; ++i) hx[i] = 2176.0 + (i % 97) * 0.5;
    for (int j = 0; j < kk; ++j) hc[j] = 1.0;
    double *dot, *xn, *cn, *out; int* flag;
    cudaMalloc(&dot, n * 8); cudaMalloc(&out, n * 8); cudaMalloc(&flag, n * 4); cudaMalloc(&xn, rows * 8); cudaMalloc(&cn, kk * 8);
    cudaMemcpy(xn, hx.data(), rows * 8, cudaMemcpyHostToDevice); cudaMemcpy(cn, hc.data(), kk * 8, cudaMemcpyHostToDevice);
    // dots: deterministic pseudo-random in [-30, 30]
    std::vector<double> hd(1 << 20);
    for (size_t i = 0; i < hd.size(); ++i) hd[i] = 60.0 * ((double)((i * 2654435761u) & 0xffffff) / 16777216.0) - 30.0;
    for (long long o = 0; o < n; o += (long long)hd.size()) cudaMemcpy(dot + o, hd.data(), (size_t)std::min<long long>(hd.size(), n - o) * 8, cudaMemcpyHostToDevice);
    int grid = 148 * per_sm;
    cudaEvent_t e0, e1; cudaEventCreate(&e0); cudaEventCreate(&e1);
    k<<<grid, threads>>>(dot, xn, cn, out, flag, n, kk, 0.25, 1);
    cudaDeviceSynchronize();
    cudaEventRecord(e0);
    k<<<grid, threads>>>(dot, xn, cn, out, flag, n, kk, 0.25, 2);
    cudaEventRecord(e1); cudaEventSynchronize(e1);
    float ms; cudaEventElapsedTime(&ms, e0, e1);
    printf("threads %d per_sm %d: %.2f ms for %lld pairs = %.2f Gpairs/s  (%s)\n", threads, per_sm, ms, n, n / ms / 1e6, cudaGetErrorString(cudaGetLastError()));
    return 0;
}
