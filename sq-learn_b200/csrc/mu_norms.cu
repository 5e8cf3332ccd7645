// K6 — the whole p-grid of the mu-norm search in two streaming passes.
//
// Replaces best_mu / linear_search / __mu (sklearn/QuantumUtility/Utility.py:
// 196-231): for every p on the grid the reference makes two full np.power
// passes, s1 = max_i sum_j |a_ij|^(2p) and s2 = max_j sum_i |a_ij|^(2(1-p)),
// with the convention that exponent 0 means count_nonzero (:198-199), and
// compares sqrt(s1*s2) with the Frobenius norm (:224).  Here the row pass and
// the column pass each read X once for ALL grid points: log2|a| is taken once
// per element and every power is one exp2.
//
// Outputs per grid point g: rowmax[g] = max over this rank's rows of the row
// sum with exponent er[g]; colsum[g, :] = column sums with exponent ec[g]
// over this rank's rows; frob2 = sum a^2.  (Across ranks: max, sum, sum.)
#include "common.cuh"
#include "sqb200.h"

namespace sq {

static const int MU_MAXP = 16;

struct MuArgs {
    const double* X;
    long long n;
    int d, np;
    double er[MU_MAXP], ec[MU_MAXP];
    unsigned long long* rowmax_bits;  // [np] max kept as raw bits (non-negative doubles order like uint64)
    double* colpart;                  // [row_blocks][np][d]
    double* frobpart;                 // [row_warps]
    long long rows_per_block;
    int row_blocks;
};

__device__ __forceinline__ double powabs(double av, double l2, bool nz, double e) {
    // |a|^e with the reference's conventions: exponent 0 counts non-zeros
    if (!nz) return 0.0;
    if (e == 0.0) return 1.0;
    if (e == 2.0) return av * av;      // (every grid has 2p = 2 or 2(1-p) = 2 at its ends, and the extra Frobenius row)
    if (e == 1.0) return av;
    return exp2(e * l2);
}

// warp per row over all d: row sums for every grid point, running max, sum of squares
__global__ void mu_row_kernel(MuArgs a) {
    const int lane = threadIdx.x & 31;
    const long long w0 = ((long long)blockIdx.x * blockDim.x + threadIdx.x) >> 5;
    const long long nw = ((long long)gridDim.x * blockDim.x) >> 5;
    double best[MU_MAXP];
#pragma unroll
    for (int g = 0; g < MU_MAXP; ++g) best[g] = 0.0;
    double frob = 0.0;
    for (long long w = w0; w < a.n; w += nw) {
        const double* x = a.X + w * (long long)a.d;
        double rs[MU_MAXP];
#pragma unroll
        for (int g = 0; g < MU_MAXP; ++g) rs[g] = 0.0;
        for (int c = lane; c < a.d; c += 32) {
            double v = x[c];
            double av = fabs(v);
            bool nz = av != 0.0;
            double l2 = nz ? log2(av) : 0.0;
            frob = fma(v, v, frob);
#pragma unroll
            for (int g = 0; g < MU_MAXP; ++g)
                if (g < a.np) rs[g] += powabs(av, l2, nz, a.er[g]);
        }
#pragma unroll
        for (int g = 0; g < MU_MAXP; ++g)
            if (g < a.np) best[g] = fmax(best[g], warp_sum(rs[g]));
    }
    frob = warp_sum(frob);
    if (lane == 0) {
        a.frobpart[w0] = frob;
#pragma unroll
        for (int g = 0; g < MU_MAXP; ++g)
            if (g < a.np) atomicMax(&a.rowmax_bits[g], (unsigned long long)__double_as_longlong(best[g]));
    }
}

// thread per column, block per (row range, 256-column chunk): column sums in registers
__global__ void __launch_bounds__(256) mu_col_kernel(MuArgs a) {
    const int c = blockIdx.y * 256 + threadIdx.x;
    const long long r0 = (long long)blockIdx.x * a.rows_per_block;
    long long r1 = r0 + a.rows_per_block;
    if (r1 > a.n) r1 = a.n;
    double cs[MU_MAXP];
#pragma unroll
    for (int g = 0; g < MU_MAXP; ++g) cs[g] = 0.0;
    if (c < a.d) {
        for (long long r = r0; r < r1; ++r) {
            double av = fabs(a.X[r * (long long)a.d + c]);
            bool nz = av != 0.0;
            double l2 = nz ? log2(av) : 0.0;
#pragma unroll
            for (int g = 0; g < MU_MAXP; ++g)
                if (g < a.np) cs[g] += powabs(av, l2, nz, a.ec[g]);
        }
#pragma unroll
        for (int g = 0; g < MU_MAXP; ++g)
            if (g < a.np) a.colpart[((long long)blockIdx.x * a.np + g) * a.d + c] = cs[g];
    }
}

__global__ void mu_combine_kernel(const double* __restrict__ colpart, const double* __restrict__ frobpart, int row_blocks,
                                  int row_warps, int np, int d, double* __restrict__ colsum, double* __restrict__ frob2) {
    long long idx = (long long)blockIdx.x * blockDim.x + threadIdx.x;
    if (idx < (long long)np * d) {
        double t = 0.0;
        for (int b = 0; b < row_blocks; ++b) t += colpart[(long long)b * np * d + idx];
        colsum[idx] = t;
    }
    if (idx == 0) {
        double t = 0.0;
        for (int b = 0; b < row_warps; ++b) t += frobpart[b];
        *frob2 = t;
    }
}

struct MuGeom { int row_blocks, col_chunks, row_grid, row_warps; long long rows_per_block; };
static MuGeom mu_geom(long long n, int d) {
    MuGeom g;
    g.col_chunks = (d + 255) / 256;
    int target = device_sm_count() * 8 / g.col_chunks;
    if (target < 1) target = 1;
    if (target > n) target = (int)n;
    g.rows_per_block = (n + target - 1) / target;
    g.row_blocks = (int)((n + g.rows_per_block - 1) / g.rows_per_block);
    g.row_grid = device_sm_count() * 8;
    g.row_warps = g.row_grid * 8;
    return g;
}

}  // namespace sq

using namespace sq;

extern "C" size_t sq_mu_norms_workspace_bytes(long long n, int d, int n_p) {
    MuGeom g = mu_geom(n, d);
    return align_up((size_t)g.row_blocks * n_p * d * sizeof(double), 256) + align_up((size_t)g.row_warps * sizeof(double), 256);
}

extern "C" int sq_mu_norms(const double* X, long long n, int d, const double* exp_row_host, const double* exp_col_host,
                           int n_p, double* out_rowmax, double* out_colsum, double* out_frob2, void* ws,
                           size_t ws_bytes, void* stream_) {
    cudaStream_t stream = (cudaStream_t)stream_;
    SQ_REQUIRE(X && exp_row_host && exp_col_host && out_rowmax && out_colsum && out_frob2, "sq_mu_norms: null pointer");
    SQ_REQUIRE(n >= 1 && n < 2147483647LL && d >= 1, "sq_mu_norms: bad shape (n=%lld rows must be below 2^31 - 1)", n);
    SQ_REQUIRE(n_p >= 1 && n_p <= MU_MAXP, "sq_mu_norms: n_p=%d outside [1,%d]", n_p, MU_MAXP);
    size_t need = sq_mu_norms_workspace_bytes(n, d, n_p);
    if (!ws || ws_bytes < need) { set_error("sq_mu_norms: workspace needs %zu bytes, got %zu", need, ws_bytes); return SQ_ERR_WORKSPACE; }
    MuGeom g = mu_geom(n, d);
    MuArgs a;
    a.X = X; a.n = n; a.d = d; a.np = n_p;
    for (int i = 0; i < n_p; ++i) {
        SQ_REQUIRE(exp_row_host[i] >= 0.0 && exp_col_host[i] >= 0.0, "sq_mu_norms: exponents must be >= 0");
        a.er[i] = exp_row_host[i]; a.ec[i] = exp_col_host[i];
    }
    a.rows_per_block = g.rows_per_block; a.row_blocks = g.row_blocks;
    a.colpart = (double*)ws;
    a.frobpart = (double*)((char*)ws + align_up((size_t)g.row_blocks * n_p * d * sizeof(double), 256));
    a.rowmax_bits = (unsigned long long*)out_rowmax;
    SQ_CUDA(cudaMemsetAsync(out_rowmax, 0, n_p * sizeof(double), stream));
    mu_row_kernel<<<g.row_grid, 256, 0, stream>>>(a);
    SQ_LAUNCH_CHECK("mu_row_kernel");
    mu_col_kernel<<<dim3(g.row_blocks, g.col_chunks), 256, 0, stream>>>(a);
    SQ_LAUNCH_CHECK("mu_col_kernel");
    long long npd = (long long)n_p * d;
    mu_combine_kernel<<<(unsigned)((npd + 255) / 256), 256, 0, stream>>>(a.colpart, a.frobpart, g.row_blocks, g.row_warps, n_p, d, out_colsum, out_frob2);
    SQ_LAUNCH_CHECK("mu_combine_kernel");
    return SQ_OK;
}
