// Microbenchmark: what DMMA.8x8x4 issue patterns reach the fp64 tensor peak on sm_100a?
// build: nvcc -gencode arch=compute_100a,code=sm_100a -O3 -o dmma_microbench dmma_microbench.cu
#include <cstdio>
#include <cuda_runtime.h>

__device__ __forceinline__ void dmma(double& d0, double& d1, double a, double b) {
    asm volatile("mma.sync.aligned.m8n8k4.row.col.f64.f64.f64.f64 {%0,%1}, {%2}, {%3}, {%0,%1};\n"
                 : "+d"(d0), "+d"(d1) : "d"(a), "d"(b));
}

// MODE 0: registers only.  MODE 1: operands from shared memory, MI+NI LDS per MI*NI DMMA.
// MODE 2: MODE 1 + __syncthreads every 4 k-steps.
template <int MI, int NI, int MODE>
__global__ void k(double* out, int iters) {
    extern __shared__ double sm[];
    const int lane = threadIdx.x & 31, warp = threadIdx.x >> 5;
    for (int i = threadIdx.x; i < 8192; i += blockDim.x) sm[i] = 1e-3 * i;
    __syncthreads();
    double acc[MI][NI][2];
#pragma unroll
    for (int i = 0; i < MI; ++i)
#pragma unroll
        for (int j = 0; j < NI; ++j) acc[i][j][0] = acc[i][j][1] = 0.0;
    double af[MI], bf[NI];
#pragma unroll
    for (int i = 0; i < MI; ++i) af[i] = lane * 0.5 + i;
#pragma unroll
    for (int j = 0; j < NI; ++j) bf[j] = lane * 0.25 + j;
    for (int it = 0; it < iters; ++it) {
#pragma unroll
        for (int kb = 0; kb < 4; ++kb) {
            if (MODE >= 1) {
                const double* p = sm + ((it + kb) & 7) * 1024 + (warp & 3) * 32 + lane;
#pragma unroll
                for (int i = 0; i < MI; ++i) af[i] = p[i * 32 * 4];
#pragma unroll
                for (int j = 0; j < NI; ++j) bf[j] = p[512 + j * 32 * 4 - (j >= 4 ? 4 * 128 - 32 : 0)];
            }
#pragma unroll
            for (int i = 0; i < MI; ++i)
#pragma unroll
                for (int j = 0; j < NI; ++j) dmma(acc[i][j][0], acc[i][j][1], af[i], bf[j]);
        }
        if (MODE == 2) __syncthreads();
    }
    double s = 0;
#pragma unroll
    for (int i = 0; i < MI; ++i)
#pragma unroll
        for (int j = 0; j < NI; ++j) s += acc[i][j][0] + acc[i][j][1];
    out[blockIdx.x * blockDim.x + threadIdx.x] = s;
}

template <int MI, int NI, int MODE>
void run(int warps, int ctas_per_sm, double* out) {
    int iters = 2000;
    int grid = 148 * ctas_per_sm;
    cudaFuncSetAttribute(k<MI, NI, MODE>, cudaFuncAttributeMaxDynamicSharedMemorySize, 65536);
    k<MI, NI, MODE><<<grid, warps * 32, 65536>>>(out, 10);
    cudaEvent_t e0, e1;
    cudaEventCreate(&e0); cudaEventCreate(&e1);
    cudaEventRecord(e0);
    k<MI, NI, MODE><<<grid, warps * 32, 65536>>>(out, iters);
    cudaEventRecord(e1);
    cudaEventSynchronize(e1);
    float ms; cudaEventElapsedTime(&ms, e0, e1);
    double flops = 2.0 * 256 * (double)MI * NI * 4 * iters * warps * grid;
    printf("MI=%d NI=%d mode=%d warps/cta=%2d ctas/sm=%d : %7.2f TFLOP/s  (%s)\n", MI, NI, MODE, warps, ctas_per_sm,
           flops / (ms * 1e-3) / 1e12, cudaGetErrorString(cudaGetLastError()));
}

int main() {
    setvbuf(stdout, NULL, _IONBF, 0);
    int ndev = 0; cudaError_t e = cudaGetDeviceCount(&ndev);
    printf("devices %d (%s)\n", ndev, cudaGetErrorString(e));
    if (e != cudaSuccess || ndev == 0) return 1;
    double* out = nullptr; e = cudaMalloc(&out, 148 * 4 * 1024 * 8);
    printf("malloc %s\n", cudaGetErrorString(e));
    if (e != cudaSuccess) return 1;
    run<4, 4, 0>(4, 1, out); run<4, 4, 0>(8, 1, out); run<4, 4, 0>(16, 1, out);
    run<2, 2, 0>(4, 1, out); run<2, 2, 0>(8, 1, out); run<2, 2, 0>(16, 1, out); run<1, 1, 0>(16, 1, out); run<1, 1, 0>(32, 1, out);
    run<4, 8, 0>(4, 1, out); run<4, 8, 0>(8, 1, out);
    run<4, 4, 1>(4, 1, out); run<4, 4, 2>(4, 1, out); run<4, 4, 1>(8, 1, out); run<4, 4, 2>(8, 1, out); run<4, 4, 1>(16, 1, out); run<4, 4, 2>(16, 1, out);
    run<4, 4, 1>(8, 2, out); run<4, 4, 2>(8, 2, out);
    run<4, 8, 1>(4, 1, out); run<4, 8, 1>(8, 1, out); run<4, 8, 2>(8, 1, out); run<8, 4, 1>(8, 1, out);
    run<2, 4, 1>(16, 1, out); run<2, 4, 2>(16, 1, out);
    return 0;
}
