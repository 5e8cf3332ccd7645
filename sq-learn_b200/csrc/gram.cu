// K4 — centred Gram matrix G = (X - mu)^T (X - mu) on the fp64 tensor cores,
// plus the column sums that give mu.
//
// Replaces the LAPACK SVD of the centred data in qPCA._fit_full
// (sklearn/decomposition/_qPCA.py:578-583) and np.linalg.svd in qMeans_.fit
// (sklearn/cluster/_dmeans.py:1244) by a new route to the same quantities:
// eigenpairs of G give sigma = sqrt(lambda) and the right singular vectors
// (the small d x d eigendecomposition runs in cuSOLVER, off the hot path).
// Centring happens when fragments are loaded (x - mu), never as G - n mu mu^T.
//
// One CTA accumulates a 128 x 128 block (bi <= bj) of G over a slice of rows
// with DMMA.8x8x4; both operands are the same row-major X tile, read
// transposed (A[m][k] = X[k][m]).  Row stride 132 doubles makes the
// fragment reads conflict free.  Slice partials are combined in a fixed
// order (deterministic) and mirrored into the lower triangle.
#include "common.cuh"
#include "sqb200.h"
#include <cstdlib>

namespace sq {

__device__ __forceinline__ double lds_f64(unsigned addr) {
    double v;
    asm volatile("ld.shared.f64 %0, [%1];" : "=d"(v) : "r"(addr));
    return v;
}

static const int GB = 128;       // block of features
static const int GR = 16;        // rows per chunk
static const int GSTRIDE = 132;  // smem row stride in doubles (== 4 mod 16 -> conflict-free fragments)
static const int GSTAGES = 4;
static const int GTHREADS = 512; // 16 warps as 4 x 4, warp tile 32 x 32

struct GramArgs {
    const double* X;
    const double* mu;   // [d] or null
    double* partial;    // [slices][pairs][GB*GB]
    long long n;
    int d, nblk, pairs;
    long long rows_per_slice;
};

__device__ __forceinline__ void gram_load(double* stage, const GramArgs& a, long long r0, long long rend, int fi, int fj,
                                          bool diag, int tid, bool vec16) {
    // stage: [2][GR][GSTRIDE]; second half unused on the diagonal
    const int halves = diag ? 1 : 2;
    if (vec16) {
        const int PPR = GB / 2;  // 16-byte pieces per row
        for (int p = tid; p < halves * GR * PPR; p += GTHREADS) {
            int h = p / (GR * PPR), q = p % (GR * PPR);
            int row = q / PPR, c = (q % PPR) * 2;
            int f = (h ? fj : fi) + c;
            long long gr = r0 + row;
            bool ok = gr < rend && f < a.d;
            const double* src = a.X + (ok ? gr * (long long)a.d + f : 0);
            cp_async16(stage + (h * GR + row) * GSTRIDE + c, src, ok);
        }
    } else {
        for (int p = tid; p < halves * GR * GB; p += GTHREADS) {
            int h = p / (GR * GB), q = p % (GR * GB);
            int row = q / GB, c = q % GB;
            int f = (h ? fj : fi) + c;
            long long gr = r0 + row;
            bool ok = gr < rend && f < a.d;
            const double* src = a.X + (ok ? gr * (long long)a.d + f : 0);
            cp_async8(stage + (h * GR + row) * GSTRIDE + c, src, ok);
        }
    }
}

__global__ void __launch_bounds__(GTHREADS, 1) gram_kernel(GramArgs a, int vec16) {
    extern __shared__ __align__(16) double smem[];
    const int tid = threadIdx.x, lane = tid & 31, warp = tid >> 5;
    const int wm = warp >> 2, wn = warp & 3;
    const int lq = lane >> 2, lr4 = lane & 3;
    // pair index -> (bi, bj), bi <= bj
    int pr = blockIdx.x, bi = 0;
    while (pr >= a.nblk - bi) { pr -= a.nblk - bi; ++bi; }
    const int bj = bi + pr;
    const bool diag = (bi == bj);
    const int fi = bi * GB, fj = bj * GB;
    const long long r0 = (long long)blockIdx.y * a.rows_per_slice;
    long long rend = r0 + a.rows_per_slice;
    if (rend > a.n) rend = a.n;
    const int chunks = (int)((rend - r0 + GR - 1) / GR);
    const int STAGE = 2 * GR * GSTRIDE;

    double mui[4], muj[4];
#pragma unroll
    for (int i = 0; i < 4; ++i) {
        int f = fi + wm * 32 + i * 8 + lq;
        mui[i] = (a.mu && f < a.d) ? a.mu[f] : 0.0;
        f = fj + wn * 32 + i * 8 + lq;
        muj[i] = (a.mu && f < a.d) ? a.mu[f] : 0.0;
    }
    double acc[4][4][2];
#pragma unroll
    for (int i = 0; i < 4; ++i)
#pragma unroll
        for (int j = 0; j < 4; ++j) acc[i][j][0] = acc[i][j][1] = 0.0;

    for (int s = 0; s < GSTAGES - 1; ++s) {
        if (s < chunks) gram_load(smem + s * STAGE, a, r0 + (long long)s * GR, rend, fi, fj, diag, tid, vec16);
        cp_async_commit();
    }
    for (int c = 0; c < chunks; ++c) {
        cp_async_wait<GSTAGES - 2>();
        __syncthreads();
        int nx = c + GSTAGES - 1;
        if (nx < chunks) gram_load(smem + (nx % GSTAGES) * STAGE, a, r0 + (long long)nx * GR, rend, fi, fj, diag, tid, vec16);
        cp_async_commit();
        const double* Ai = smem + (c % GSTAGES) * STAGE;
        const double* Aj = diag ? Ai : Ai + GR * GSTRIDE;
        const long long rbase = r0 + (long long)c * GR;
#pragma unroll
        for (int kb = 0; kb < GR / 4; ++kb) {
            const int row = kb * 4 + lr4;
            // rows past the end were zero-filled; centring must not turn them into -mu
            const bool live = (rbase + row) < rend;
            double af[4], bf[4];
#pragma unroll
            for (int i = 0; i < 4; ++i) {
                double v = Ai[row * GSTRIDE + wm * 32 + i * 8 + lq];
                int f = fi + wm * 32 + i * 8 + lq;
                af[i] = (live && f < a.d) ? v - mui[i] : 0.0;
            }
#pragma unroll
            for (int j = 0; j < 4; ++j) {
                double v = Aj[row * GSTRIDE + wn * 32 + j * 8 + lq];
                int f = fj + wn * 32 + j * 8 + lq;
                bf[j] = (live && f < a.d) ? v - muj[j] : 0.0;
            }
#pragma unroll
            for (int i = 0; i < 4; ++i)
#pragma unroll
                for (int j = 0; j < 4; ++j) dmma884(acc[i][j][0], acc[i][j][1], af[i], bf[j]);
        }
    }
    cp_async_wait<0>();
    double* out = a.partial + ((long long)blockIdx.y * a.pairs + blockIdx.x) * (GB * GB);
#pragma unroll
    for (int i = 0; i < 4; ++i)
#pragma unroll
        for (int j = 0; j < 4; ++j) {
            int r = wm * 32 + i * 8 + lq, cc = wn * 32 + j * 8 + lr4 * 2;
            *reinterpret_cast<double2*>(out + r * GB + cc) = make_double2(acc[i][j][0], acc[i][j][1]);
        }
}

// ---- TMA variant (no centring; d even, X 16-byte aligned) ------------------------------------
// The X tile arrives as boxes {4 features, GT_R rows}: slab layout [feature/4][row][4], so the
// (8 features x 4 rows) DMMA fragment of either operand is two runs of 128 contiguous bytes and
// needs no padding.  One elected thread issues the boxes; completion on an mbarrier per stage.
static const int GT_R = 32;        // rows per chunk
static const int GT_STAGES = 3;
__global__ void __launch_bounds__(GTHREADS, 1) gram_tma_kernel(GramArgs a, const __grid_constant__ CUtensorMap tm) {
    extern __shared__ __align__(1024) double smem[];
    const int tid = threadIdx.x, lane = tid & 31, warp = tid >> 5;
    const int wm = warp >> 2, wn = warp & 3;
    const int lq = lane >> 2, lr4 = lane & 3;
    int pr = blockIdx.x, bi = 0;
    while (pr >= a.nblk - bi) { pr -= a.nblk - bi; ++bi; }
    const int bj = bi + pr;
    const bool diag = (bi == bj);
    const int fi = bi * GB, fj = bj * GB;
    const long long r0 = (long long)blockIdx.y * a.rows_per_slice;
    long long rend = r0 + a.rows_per_slice;
    if (rend > a.n) rend = a.n;
    const int chunks = (int)((rend - r0 + GT_R - 1) / GT_R);
    constexpr int HALF = GB * GT_R;            // doubles of one operand's tile
    constexpr int STAGE = 2 * HALF;
    unsigned long long* full_bar = (unsigned long long*)(smem + GT_STAGES * STAGE);
    if (tid == 0) {
        for (int s = 0; s < GT_STAGES; ++s) mbar_init(&full_bar[s], 1);
        asm volatile("fence.mbarrier_init.release.cluster;" ::: "memory");
        asm volatile("fence.proxy.async.shared::cta;" ::: "memory");
    }
    __syncthreads();
    // rows past rend inside the last chunk of a slice belong to the next slice: they must not be counted
    // twice, so the slice length is a multiple of GT_R (host) and only the matrix end is ragged (zero fill).
    auto issue = [&](int c) {
        if (tid == 0 && c < chunks) {
            double* st = smem + (c % GT_STAGES) * STAGE;
            unsigned long long* bar = &full_bar[c % GT_STAGES];
            mbar_expect_tx(bar, (diag ? HALF : STAGE) * 8);
            const int row = (int)(r0 + (long long)c * GT_R);
#pragma unroll 4
            for (int g = 0; g < GB / 4; ++g) {
                tma_load_2d(st + g * GT_R * 4, &tm, fi + g * 4, row, bar);
                if (!diag) tma_load_2d(st + HALF + g * GT_R * 4, &tm, fj + g * 4, row, bar);
            }
        }
    };
    for (int s = 0; s < GT_STAGES - 1; ++s) issue(s);

    double acc[4][4][2];
#pragma unroll
    for (int i = 0; i < 4; ++i)
#pragma unroll
        for (int j = 0; j < 4; ++j) acc[i][j][0] = acc[i][j][1] = 0.0;
    const unsigned smem_u32 = (unsigned)__cvta_generic_to_shared(smem);
    // fragment base: feature group (w*8 + i*2 + lq/4), row lr4, sub-feature lq%4
    const unsigned fa = (unsigned)((((wm * 8 + (lq >> 2)) * GT_R + lr4) * 4 + (lq & 3)) * 8);
    const unsigned fb = (unsigned)((((wn * 8 + (lq >> 2)) * GT_R + lr4) * 4 + (lq & 3)) * 8);
    unsigned phase = 0;
    for (int c = 0; c < chunks; ++c) {
        const int stg = c % GT_STAGES;
        mbar_wait(&full_bar[stg], phase);
        if (stg == GT_STAGES - 1) phase ^= 1;
        __syncthreads();                       // everyone is done with the stage refilled below
        issue(c + GT_STAGES - 1);
        const unsigned sa = smem_u32 + (unsigned)(stg * STAGE * 8) + fa;
        const unsigned sb = smem_u32 + (unsigned)(stg * STAGE * 8) + (diag ? 0u : (unsigned)(HALF * 8)) + fb;
#pragma unroll
        for (int kb = 0; kb < GT_R / 4; ++kb) {
            double af[4], bf[4];
#pragma unroll
            for (int i = 0; i < 4; ++i) af[i] = lds_f64(sa + (unsigned)(((i * 2 * GT_R) + kb * 4) * 32));
#pragma unroll
            for (int j = 0; j < 4; ++j) bf[j] = lds_f64(sb + (unsigned)(((j * 2 * GT_R) + kb * 4) * 32));
#pragma unroll
            for (int i = 0; i < 4; ++i)
#pragma unroll
                for (int j = 0; j < 4; ++j) dmma884(acc[i][j][0], acc[i][j][1], af[i], bf[j]);
        }
    }
    double* out = a.partial + ((long long)blockIdx.y * a.pairs + blockIdx.x) * (GB * GB);
#pragma unroll
    for (int i = 0; i < 4; ++i)
#pragma unroll
        for (int j = 0; j < 4; ++j) {
            int r = wm * 32 + i * 8 + lq, cc = wn * 32 + j * 8 + lr4 * 2;
            *reinterpret_cast<double2*>(out + r * GB + cc) = make_double2(acc[i][j][0], acc[i][j][1]);
        }
}

__global__ void gram_combine_kernel(const double* __restrict__ partial, int slices, int pairs, int nblk, int d,
                                    double* __restrict__ G) {
    long long idx = (long long)blockIdx.x * blockDim.x + threadIdx.x;
    if (idx >= (long long)d * d) return;
    int r = (int)(idx / d), c = (int)(idx % d);
    int rr = r <= c ? r : c, cc = r <= c ? c : r;  // upper-triangular source
    int bi = rr / GB, bj = cc / GB;
    int li = rr % GB, lj = cc % GB;
    if (bi == bj && li > lj) { int t = li; li = lj; lj = t; }  // inside a diagonal block both halves are valid
    int pair = 0;
    for (int b = 0; b < bi; ++b) pair += nblk - b;
    pair += bj - bi;
    double t = 0.0;
    for (int s = 0; s < slices; ++s) t += partial[((long long)s * pairs + pair) * (GB * GB) + li * GB + lj];
    G[idx] = t;
}

// column sums in two deterministic stages
__global__ void colsum_partial_kernel(const double* __restrict__ X, long long n, int d, long long rows_per_block,
                                      double* __restrict__ part) {
    const int c = blockIdx.y * blockDim.x + threadIdx.x;
    if (c >= d) return;
    const long long r0 = (long long)blockIdx.x * rows_per_block;
    long long r1 = r0 + rows_per_block;
    if (r1 > n) r1 = n;
    double s0 = 0.0, s1 = 0.0, s2 = 0.0, s3 = 0.0;
    long long r = r0;
    for (; r + 3 < r1; r += 4) {
        s0 += X[r * (long long)d + c];
        s1 += X[(r + 1) * (long long)d + c];
        s2 += X[(r + 2) * (long long)d + c];
        s3 += X[(r + 3) * (long long)d + c];
    }
    for (; r < r1; ++r) s0 += X[r * (long long)d + c];
    part[(long long)blockIdx.x * d + c] = (s0 + s1) + (s2 + s3);
}
__global__ void colsum_final_kernel(const double* __restrict__ part, int blocks, int d, double* __restrict__ out) {
    const int c = blockIdx.x * blockDim.x + threadIdx.x;
    if (c >= d) return;
    double t = 0.0;
    for (int b = 0; b < blocks; ++b) t += part[(long long)b * d + c];
    out[c] = t;
}

// Y = X - mu (row broadcast), in place allowed
__global__ void sub_rowvec_kernel(const double* __restrict__ X, const double* __restrict__ mu, long long total, int d,
                                  double* __restrict__ Y) {
    for (long long i = (long long)blockIdx.x * blockDim.x + threadIdx.x; i < total; i += (long long)gridDim.x * blockDim.x)
        Y[i] = X[i] - mu[i % d];
}

struct GramGeom { int nblk, pairs, slices; long long rows_per_slice; };
static GramGeom gram_geom(long long n, int d) {
    GramGeom g;
    g.nblk = (d + GB - 1) / GB;
    g.pairs = g.nblk * (g.nblk + 1) / 2;
    int target = (device_sm_count() * 2 + g.pairs - 1) / g.pairs;
    long long maxs = (n + 4 * GR - 1) / (4 * GR);
    if (target > maxs) target = (int)maxs;
    if (target < 1) target = 1;
    g.rows_per_slice = ((n + target - 1) / target + 31) / 32 * 32;   // multiple of both chunk sizes (16, 32)
    g.slices = (int)((n + g.rows_per_slice - 1) / g.rows_per_slice);
    return g;
}
static int colsum_blocks(long long n, int d) {
    int cchunks = (d + 127) / 128;
    int b = device_sm_count() * 8 / cchunks;
    if (b < 1) b = 1;
    if (b > n) b = (int)n;
    return b;
}

}  // namespace sq

using namespace sq;

extern "C" size_t sq_gram_workspace_bytes(long long n, int d) {
    GramGeom g = gram_geom(n, d);
    return (size_t)g.slices * g.pairs * GB * GB * sizeof(double);
}

extern "C" int sq_gram_centered(const double* X, long long n, int d, const double* mean, double* out_gram, void* ws,
                                size_t ws_bytes, void* stream_) {
    cudaStream_t stream = (cudaStream_t)stream_;
    SQ_REQUIRE(X && out_gram && n >= 1 && d >= 1, "sq_gram_centered: bad arguments");
    SQ_REQUIRE(n < 2147483647LL, "sq_gram_centered: n=%lld rows exceed the 2^31 - 1 the row coordinates hold", n);
    size_t need = sq_gram_workspace_bytes(n, d);
    if (!ws || ws_bytes < need) { set_error("sq_gram_centered: workspace needs %zu bytes, got %zu", need, ws_bytes); return SQ_ERR_WORKSPACE; }
    GramGeom g = gram_geom(n, d);
    GramArgs a;
    a.X = X; a.mu = mean; a.partial = (double*)ws; a.n = n; a.d = d; a.nblk = g.nblk; a.pairs = g.pairs;
    a.rows_per_slice = g.rows_per_slice;
    size_t smem = (size_t)GSTAGES * 2 * GR * GSTRIDE * sizeof(double);
    SQ_CUDA(cudaFuncSetAttribute(gram_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem));
    int vec16 = ((d & 1) == 0 && ((uintptr_t)X & 15) == 0) ? 1 : 0;
    CUtensorMap tm;
    static const bool no_tma = getenv("SQ_GRAM_NO_TMA") != nullptr;   // A/B timing switch, read once
    bool use_tma = vec16 && !mean && !no_tma &&
                   make_tensor_map_2d(&tm, X, (unsigned long long)d, (unsigned long long)n, 4, GT_R) == 0;
    if (use_tma) {
        size_t sm2 = (size_t)GT_STAGES * 2 * GB * GT_R * sizeof(double) + GT_STAGES * sizeof(unsigned long long);
        SQ_CUDA(cudaFuncSetAttribute(gram_tma_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)sm2));
        gram_tma_kernel<<<dim3(g.pairs, g.slices), GTHREADS, sm2, stream>>>(a, tm);
        SQ_LAUNCH_CHECK("gram_tma_kernel");
    } else {
        gram_kernel<<<dim3(g.pairs, g.slices), GTHREADS, smem, stream>>>(a, vec16);
        SQ_LAUNCH_CHECK("gram_kernel");
    }
    long long dd = (long long)d * d;
    gram_combine_kernel<<<(unsigned)((dd + 255) / 256), 256, 0, stream>>>(a.partial, g.slices, g.pairs, g.nblk, d, out_gram);
    SQ_LAUNCH_CHECK("gram_combine_kernel");
    return SQ_OK;
}

extern "C" size_t sq_col_sums_workspace_bytes(long long n, int d) { return (size_t)colsum_blocks(n, d) * d * sizeof(double); }

extern "C" int sq_col_sums(const double* X, long long n, int d, double* out, void* ws, size_t ws_bytes, void* stream_) {
    cudaStream_t stream = (cudaStream_t)stream_;
    SQ_REQUIRE(X && out && n >= 1 && d >= 1, "sq_col_sums: bad arguments");
    SQ_REQUIRE(n < 2147483647LL, "sq_col_sums: n=%lld rows exceed 2^31 - 1", n);
    size_t need = sq_col_sums_workspace_bytes(n, d);
    if (!ws || ws_bytes < need) { set_error("sq_col_sums: workspace needs %zu bytes, got %zu", need, ws_bytes); return SQ_ERR_WORKSPACE; }
    int blocks = colsum_blocks(n, d);
    long long rpb = (n + blocks - 1) / blocks;
    blocks = (int)((n + rpb - 1) / rpb);
    colsum_partial_kernel<<<dim3(blocks, (d + 127) / 128), 128, 0, stream>>>(X, n, d, rpb, (double*)ws);
    SQ_LAUNCH_CHECK("colsum_partial_kernel");
    colsum_final_kernel<<<(d + 127) / 128, 128, 0, stream>>>((double*)ws, blocks, d, out);
    SQ_LAUNCH_CHECK("colsum_final_kernel");
    return SQ_OK;
}

extern "C" int sq_sub_rowvec(const double* X, const double* mu, long long n, int d, double* Y, void* stream_) {
    cudaStream_t stream = (cudaStream_t)stream_;
    SQ_REQUIRE(X && mu && Y && n >= 0 && d >= 1, "sq_sub_rowvec: bad arguments");
    if (n == 0) return SQ_OK;
    sub_rowvec_kernel<<<device_sm_count() * 16, 256, 0, stream>>>(X, mu, n * (long long)d, d, Y);
    SQ_LAUNCH_CHECK("sub_rowvec_kernel");
    return SQ_OK;
}
