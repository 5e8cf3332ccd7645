// K3 — q-means M-step: per-cluster sums and counts, deterministic.
//
// Replaces _centers_update (sklearn/cluster/_dmeans.py:814-823): for every
// label, np.sum of the member rows divided by the member count.  The
// reference scans all n labels once per cluster in interpreted Python
// (O(n*k)); here one pass over X suffices.
//
// A k x d fp64 accumulator table does not fit in shared memory at the headline
// shape (256 x 128 x 8 B = 256 KB), and fp64 atomics would make the sums
// depend on scheduling.  Instead the rows are bucketed by label first
// (stable counting sort of row indices: histogram -> scan -> scatter, all on
// 4-byte labels), then each bucket is reduced in fixed-size items whose
// partial sums are combined in a fixed order.  X itself is read exactly once,
// in label order; rows are contiguous d*8-byte records so the gather stays
// coalesced.  Algorithmic traffic: 8*n*d bytes of X + ~16*n bytes of labels
// and row indices.
#include "common.cuh"
#include "sqb200.h"

namespace sq {

static const int CW_WARPS_PER_BLOCK = 8;
static const int FCHUNK = 512;       // features handled per pass of an item

struct CentroidWs {
    int* hist;        // [k][W] per-warp label counts, then exclusive offsets
    int* perm;        // [n]   row indices grouped by label
    int* base;        // [k+1] first slot of each label in perm
    int* item_off;    // [k+1] first item of each label
    int* scan_done;   // [1]   blocks of label_scan_kernel that have finished (reset by the last one)
    double* partial;  // [max_items][d]
};

struct CentroidGeom {
    int W;             // number of row slices (one warp each)
    long long rows_per_warp;
    int blocks;
    long long max_items;
    int item_rows;     // rows per reduction item
};

static CentroidGeom centroid_geom(long long n, int k) {
    CentroidGeom g;
    int blocks = device_sm_count() * 4;
    long long maxblocks = (n + 32LL * CW_WARPS_PER_BLOCK - 1) / (32LL * CW_WARPS_PER_BLOCK);
    if (blocks > maxblocks) blocks = (int)maxblocks;
    if (blocks < 1) blocks = 1;
    g.blocks = blocks;
    g.W = blocks * CW_WARPS_PER_BLOCK;
    long long rpw = (n + g.W - 1) / g.W;
    g.rows_per_warp = (rpw + 31) / 32 * 32;
    // enough items to fill the machine several times over, at most 2048 rows each
    long long ir = n / ((long long)device_sm_count() * 8);
    int item_rows = 64;
    while (item_rows < ir && item_rows < 2048) item_rows <<= 1;
    g.item_rows = item_rows;
    g.max_items = n / item_rows + k + 1;
    return g;
}

static size_t centroid_ws_layout(long long n, int d, int k, CentroidWs* w, char* base) {
    CentroidGeom g = centroid_geom(n, k);
    size_t off = 0;
    auto take = [&](size_t bytes) { size_t o = off; off = align_up(off + bytes, 256); return base ? base + o : nullptr; };
    char* p;
    p = take((size_t)k * g.W * sizeof(int));          if (w) w->hist = (int*)p;
    p = take((size_t)n * sizeof(int));                if (w) w->perm = (int*)p;
    p = take((size_t)(k + 1) * sizeof(int));          if (w) w->base = (int*)p;
    p = take((size_t)(k + 1) * sizeof(int));          if (w) w->item_off = (int*)p;
    p = take(256);                                    if (w) w->scan_done = (int*)p;
    p = take((size_t)g.max_items * d * sizeof(double)); if (w) w->partial = (double*)p;
    return off;
}

// ---- 1. per-warp label histogram -------------------------------------------
__global__ void label_hist_kernel(const int* __restrict__ labels, long long n, int k, long long rows_per_warp, int W,
                                  int* __restrict__ hist, int* __restrict__ bad, int* __restrict__ scan_done) {
    extern __shared__ int sh[];
    const int lane = threadIdx.x & 31, wl = threadIdx.x >> 5;
    if (blockIdx.x == 0 && threadIdx.x == 0) *scan_done = 0;     // counter of the next kernel (workspace is uninitialised)
    const int w = blockIdx.x * CW_WARPS_PER_BLOCK + wl;
    int* h = sh + wl * k;
    for (int j = lane; j < k; j += 32) h[j] = 0;
    __syncwarp();
    long long r0 = (long long)w * rows_per_warp, r1 = r0 + rows_per_warp;
    if (r1 > n) r1 = n;
    for (long long r = r0 + lane; r < r0 + rows_per_warp; r += 32) {
        int lab = (r < r1) ? labels[r] : -1;
        if (lab >= k || (lab < 0 && r < r1)) { *bad = 1; lab = -1; }
        unsigned m = __match_any_sync(0xffffffffu, lab);
        if (lab >= 0 && (m & ((1u << lane) - 1)) == 0) h[lab] += __popc(m);
        __syncwarp();
    }
    for (int j = lane; j < k; j += 32) hist[(long long)j * W + w] = h[j];
}

// ---- 2. + 3. exclusive scan of every label's row over the W slices; the block that finishes last then
// turns the k totals into label bases and item offsets (two exclusive scans over k) ----------------
__device__ __forceinline__ void label_base_scan(const int* totals, int k, int* base, int* item_off, double* counts,
                                                int ITEM_ROWS) {
    __shared__ int wsum_c[32], wsum_i[32];
    __shared__ int carry_c, carry_i;
    const int tid = threadIdx.x, lane = tid & 31, wl = tid >> 5;
    if (tid == 0) { carry_c = 0; carry_i = 0; }
    __syncthreads();
    for (int j0 = 0; j0 < k; j0 += blockDim.x) {
        const int j = j0 + tid;
        const int c = (j < k) ? totals[j] : 0;          // read before base[j] is overwritten (totals may alias base)
        const int it = (c + ITEM_ROWS - 1) / ITEM_ROWS;
        int xc = c, xi = it;
#pragma unroll
        for (int o = 1; o < 32; o <<= 1) {
            int yc = __shfl_up_sync(0xffffffffu, xc, o), yi = __shfl_up_sync(0xffffffffu, xi, o);
            if (lane >= o) { xc += yc; xi += yi; }
        }
        if (lane == 31) { wsum_c[wl] = xc; wsum_i[wl] = xi; }
        __syncthreads();
        int oc = carry_c, oi = carry_i;
        for (int w = 0; w < wl; ++w) { oc += wsum_c[w]; oi += wsum_i[w]; }
        if (j < k) {
            base[j] = oc + xc - c;
            item_off[j] = oi + xi - it;
            if (counts) counts[j] = (double)c;
        }
        __syncthreads();
        if (tid == blockDim.x - 1) { carry_c = oc + xc; carry_i = oi + xi; }
        __syncthreads();
    }
    if (tid == 0) { base[k] = carry_c; item_off[k] = carry_i; }
}

__global__ void label_scan_kernel(int* __restrict__ hist, int W, int* totals, int k, int* item_off, double* counts,
                                  int ITEM_ROWS, int* done) {
    __shared__ int wsum[32];
    __shared__ int carry_s;
    __shared__ int s_last;
    const int j = blockIdx.x, tid = threadIdx.x, lane = tid & 31, wl = tid >> 5;
    int* row = hist + (long long)j * W;
    if (tid == 0) carry_s = 0;
    __syncthreads();
    for (int base = 0; base < W; base += blockDim.x) {
        int i = base + tid;
        int v = (i < W) ? row[i] : 0;
        int x = v;
#pragma unroll
        for (int o = 1; o < 32; o <<= 1) { int y = __shfl_up_sync(0xffffffffu, x, o); if (lane >= o) x += y; }
        if (lane == 31) wsum[wl] = x;
        __syncthreads();
        if (wl == 0) {
            int s = (lane < (int)(blockDim.x >> 5)) ? wsum[lane] : 0;
#pragma unroll
            for (int o = 1; o < 32; o <<= 1) { int y = __shfl_up_sync(0xffffffffu, s, o); if (lane >= o) s += y; }
            wsum[lane] = s;  // inclusive over warps
        }
        __syncthreads();
        int carry = carry_s;
        int excl = carry + (wl ? wsum[wl - 1] : 0) + x - v;
        if (i < W) row[i] = excl;
        __syncthreads();
        if (tid == 0) carry_s = carry + wsum[(blockDim.x >> 5) - 1];
        __syncthreads();
    }
    if (tid == 0) {
        totals[j] = carry_s;
        __threadfence();                                   // the total is visible before the block counts as done
        s_last = (atomicAdd(done, 1) == (int)gridDim.x - 1);
    }
    __syncthreads();
    if (!s_last) return;
    __threadfence();
    if (tid == 0) *done = 0;                               // ready for the next call
    // totals[] aliases base[]: label_base_scan reads totals[j] before it writes base[j]
    label_base_scan(totals, k, totals, item_off, counts, ITEM_ROWS);
}

// ---- 4. stable scatter of row indices into label buckets ---------------------
__global__ void label_scatter_kernel(const int* __restrict__ labels, long long n, int k, long long rows_per_warp, int W,
                                     const int* __restrict__ hist, const int* __restrict__ base, int* __restrict__ perm) {
    extern __shared__ int sh[];
    const int lane = threadIdx.x & 31, wl = threadIdx.x >> 5;
    const int w = blockIdx.x * CW_WARPS_PER_BLOCK + wl;
    int* cur = sh + wl * k;
    for (int j = lane; j < k; j += 32) cur[j] = base[j] + hist[(long long)j * W + w];
    __syncwarp();
    long long r0 = (long long)w * rows_per_warp, r1 = r0 + rows_per_warp;
    if (r1 > n) r1 = n;
    for (long long r = r0 + lane; r < r0 + rows_per_warp; r += 32) {
        int lab = (r < r1) ? labels[r] : -1;
        if (lab >= k) lab = -1;
        unsigned m = __match_any_sync(0xffffffffu, lab);
        int rank = __popc(m & ((1u << lane) - 1));
        if (lab >= 0) perm[cur[lab] + rank] = (int)r;
        __syncwarp();
        if (lab >= 0 && rank == 0) cur[lab] += __popc(m);
        __syncwarp();
    }
}

// ---- 5. per-item partial sums -------------------------------------------------
// item -> (label, row range) by binary search in item_off; one CTA per item.
__global__ void __launch_bounds__(256) segment_sum_kernel(const double* __restrict__ X, int d, int k,
                                                          const int* __restrict__ perm, const int* __restrict__ base,
                                                          const int* __restrict__ item_off, double* __restrict__ partial,
                                                          int ITEM_ROWS) {
    __shared__ double red[CW_WARPS_PER_BLOCK][FCHUNK];
    const int lane = threadIdx.x & 31, wl = threadIdx.x >> 5;
    const int total_items = item_off[k];
    for (int item = blockIdx.x; item < total_items; item += gridDim.x) {
        int lo = 0, hi = k;  // largest j with item_off[j] <= item
        while (hi - lo > 1) { int mid = (lo + hi) >> 1; if (item_off[mid] <= item) lo = mid; else hi = mid; }
        const int j = lo;
        const int s0 = base[j] + (item - item_off[j]) * ITEM_ROWS;
        int s1 = s0 + ITEM_ROWS;
        if (s1 > base[j + 1]) s1 = base[j + 1];
        for (int f0 = 0; f0 < d; f0 += FCHUNK) {
            const int fw = (d - f0 < FCHUNK) ? d - f0 : FCHUNK;
            double acc[FCHUNK / 32];
#pragma unroll
            for (int t = 0; t < FCHUNK / 32; ++t) acc[t] = 0.0;
            if ((d & 1) == 0 && (((uintptr_t)X) & 15) == 0) {
                // lane covers feature pairs (2*lane, 2*lane+1) + 64*t
                int s = s0 + wl;
                for (; s + 3 * CW_WARPS_PER_BLOCK < s1; s += 4 * CW_WARPS_PER_BLOCK) {
                    const double* x0 = X + (long long)perm[s] * d + f0;
                    const double* x1 = X + (long long)perm[s + CW_WARPS_PER_BLOCK] * d + f0;
                    const double* x2 = X + (long long)perm[s + 2 * CW_WARPS_PER_BLOCK] * d + f0;
                    const double* x3 = X + (long long)perm[s + 3 * CW_WARPS_PER_BLOCK] * d + f0;
#pragma unroll
                    for (int t = 0; t < FCHUNK / 64; ++t) {
                        int f = 2 * lane + 64 * t;
                        if (f < fw) {
                            double2 v0 = __ldg(reinterpret_cast<const double2*>(x0 + f));
                            double2 v1 = __ldg(reinterpret_cast<const double2*>(x1 + f));
                            double2 v2 = __ldg(reinterpret_cast<const double2*>(x2 + f));
                            double2 v3 = __ldg(reinterpret_cast<const double2*>(x3 + f));
                            acc[2 * t] += (v0.x + v1.x) + (v2.x + v3.x);
                            acc[2 * t + 1] += (v0.y + v1.y) + (v2.y + v3.y);
                        }
                    }
                }
                for (; s < s1; s += CW_WARPS_PER_BLOCK) {
                    const double* x0 = X + (long long)perm[s] * d + f0;
#pragma unroll
                    for (int t = 0; t < FCHUNK / 64; ++t) {
                        int f = 2 * lane + 64 * t;
                        if (f < fw) {
                            double2 v0 = __ldg(reinterpret_cast<const double2*>(x0 + f));
                            acc[2 * t] += v0.x;
                            acc[2 * t + 1] += v0.y;
                        }
                    }
                }
                __syncthreads();  // red[] of the previous chunk has been consumed
#pragma unroll
                for (int t = 0; t < FCHUNK / 64; ++t) {
                    int f = 2 * lane + 64 * t;
                    red[wl][f] = acc[2 * t];
                    red[wl][f + 1] = acc[2 * t + 1];
                }
            } else {
                // odd d (or unaligned X): lane covers features lane + 32*t
                for (int s = s0 + wl; s < s1; s += CW_WARPS_PER_BLOCK) {
                    const double* x0 = X + (long long)perm[s] * d + f0;
#pragma unroll
                    for (int t = 0; t < FCHUNK / 32; ++t) {
                        int f = lane + 32 * t;
                        if (f < fw) acc[t] += __ldg(x0 + f);
                    }
                }
                __syncthreads();
#pragma unroll
                for (int t = 0; t < FCHUNK / 32; ++t) red[wl][lane + 32 * t] = acc[t];
            }
            __syncthreads();
            for (int f = threadIdx.x; f < fw; f += blockDim.x) {
                double t = 0.0;
#pragma unroll
                for (int w = 0; w < CW_WARPS_PER_BLOCK; ++w) t += red[w][f];
                partial[(long long)item * d + f0 + f] = t;
            }
        }
    }
}

// ---- 6. combine the items of each label in order -----------------------------
__global__ void segment_combine_kernel(const double* __restrict__ partial, const int* __restrict__ item_off, int k, int d,
                                       double* __restrict__ sums) {
    long long idx = (long long)blockIdx.x * blockDim.x + threadIdx.x;
    if (idx >= (long long)k * d) return;
    int j = (int)(idx / d), f = (int)(idx % d);
    double t = 0.0;
    for (int it = item_off[j]; it < item_off[j + 1]; ++it) t += partial[(long long)it * d + f];
    sums[idx] = t;
}

// centres = sums / counts; an empty cluster keeps its previous centre
__global__ void centroid_finalize_kernel(const double* __restrict__ sums, const double* __restrict__ counts,
                                         const double* __restrict__ c_old, int k, int d, double* __restrict__ c_new) {
    long long idx = (long long)blockIdx.x * blockDim.x + threadIdx.x;
    if (idx >= (long long)k * d) return;
    int j = (int)(idx / d);
    double c = counts[j];
    c_new[idx] = (c > 0.0) ? sums[idx] / c : (c_old ? c_old[idx] : 0.0);
}

__global__ void sqdiff_kernel(const double* __restrict__ a, const double* __restrict__ b, long long n, double* __restrict__ out) {
    // single block, fixed order: squared Frobenius distance (squared_norm(centers_old - centers), _dmeans.py:651)
    __shared__ double red[32];
    double s = 0.0;
    for (long long i = threadIdx.x; i < n; i += blockDim.x) { double df = a[i] - b[i]; s += df * df; }
    s = warp_sum(s);
    if ((threadIdx.x & 31) == 0) red[threadIdx.x >> 5] = s;
    __syncthreads();
    if (threadIdx.x == 0) {
        double t = 0.0;
        for (int w = 0; w < (int)(blockDim.x >> 5); ++w) t += red[w];
        *out = t;
    }
}

}  // namespace sq

using namespace sq;

extern "C" size_t sq_centroid_workspace_bytes(long long n, int d, int k) {
    return centroid_ws_layout(n, d, k, nullptr, nullptr);
}

extern "C" int sq_centroid_accumulate(const double* X, const int* labels, long long n, int d, int k, double* sums,
                                      double* counts, int* bad_label_flag, void* ws, size_t ws_bytes,
                                      void* stream_) {
    cudaStream_t stream = (cudaStream_t)stream_;
    SQ_REQUIRE(X && labels && sums && counts, "sq_centroid_accumulate: null pointer");
    SQ_REQUIRE(n >= 1 && n < 2147483647LL && d >= 1 && k >= 1, "sq_centroid_accumulate: bad shape n=%lld d=%d k=%d", n, d, k);
    if (k > 4096) { set_error("sq_centroid_accumulate: k=%d exceeds the supported 4096", k); return SQ_ERR_UNSUPPORTED; }
    size_t need = centroid_ws_layout(n, d, k, nullptr, nullptr);
    if (!ws || ws_bytes < need) { set_error("sq_centroid_accumulate: workspace needs %zu bytes, got %zu", need, ws_bytes); return SQ_ERR_WORKSPACE; }
    CentroidWs w;
    centroid_ws_layout(n, d, k, &w, (char*)ws);
    CentroidGeom g = centroid_geom(n, k);
    size_t sh = (size_t)CW_WARPS_PER_BLOCK * k * sizeof(int);
    if (sh > 48 * 1024) {
        SQ_CUDA(cudaFuncSetAttribute(label_hist_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)sh));
        SQ_CUDA(cudaFuncSetAttribute(label_scatter_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)sh));
    }
    int* bad = bad_label_flag ? bad_label_flag : w.item_off;  // scratch int when the caller does not care
    label_hist_kernel<<<g.blocks, 32 * CW_WARPS_PER_BLOCK, sh, stream>>>(labels, n, k, g.rows_per_warp, g.W, w.hist, bad,
                                                                         w.scan_done);
    SQ_LAUNCH_CHECK("label_hist_kernel");
    label_scan_kernel<<<k, 256, 0, stream>>>(w.hist, g.W, w.base, k, w.item_off, counts, g.item_rows, w.scan_done);
    SQ_LAUNCH_CHECK("label_scan_kernel");
    label_scatter_kernel<<<g.blocks, 32 * CW_WARPS_PER_BLOCK, sh, stream>>>(labels, n, k, g.rows_per_warp, g.W, w.hist, w.base, w.perm);
    SQ_LAUNCH_CHECK("label_scatter_kernel");
    int items_grid = (int)(g.max_items < 65535LL * 16 ? g.max_items : 65535LL * 16);
    segment_sum_kernel<<<items_grid, 256, 0, stream>>>(X, d, k, w.perm, w.base, w.item_off, w.partial, g.item_rows);
    SQ_LAUNCH_CHECK("segment_sum_kernel");
    long long kd = (long long)k * d;
    segment_combine_kernel<<<(unsigned)((kd + 255) / 256), 256, 0, stream>>>(w.partial, w.item_off, k, d, sums);
    SQ_LAUNCH_CHECK("segment_combine_kernel");
    return SQ_OK;
}

extern "C" int sq_centroid_finalize(const double* sums, const double* counts, const double* c_old, int k, int d,
                                    double* c_new, void* stream_) {
    cudaStream_t stream = (cudaStream_t)stream_;
    SQ_REQUIRE(sums && counts && c_new && k >= 1 && d >= 1, "sq_centroid_finalize: bad arguments");
    long long kd = (long long)k * d;
    centroid_finalize_kernel<<<(unsigned)((kd + 255) / 256), 256, 0, stream>>>(sums, counts, c_old, k, d, c_new);
    SQ_LAUNCH_CHECK("centroid_finalize_kernel");
    return SQ_OK;
}

extern "C" int sq_sqdiff_sum(const double* a, const double* b, long long n, double* out, void* stream_) {
    cudaStream_t stream = (cudaStream_t)stream_;
    SQ_REQUIRE(a && b && out && n >= 0, "sq_sqdiff_sum: bad arguments");
    sqdiff_kernel<<<1, 1024, 0, stream>>>(a, b, n, out);
    SQ_LAUNCH_CHECK("sqdiff_kernel");
    return SQ_OK;
}

// ---------------------------------------------------------------------------
// k-means++ step, second half: with dots[i, t] = x_i . c_t for the T candidates of the step (from sq_project),
// d2[t, i] = min(closest[i], max(|x_i|^2 + |c_t|^2 - 2 dots[i, t], 0)) - the candidate's new closest-distance row.
namespace sq {
__global__ void kpp_update_kernel(const double* __restrict__ dots, const double* __restrict__ xnorm,
                                  const double* __restrict__ cnorm, const double* __restrict__ closest, long long n,
                                  int T, double* __restrict__ d2) {
    for (long long i = (long long)blockIdx.x * blockDim.x + threadIdx.x; i < n; i += (long long)gridDim.x * blockDim.x) {
        const double xn = xnorm[i], cl = closest ? closest[i] : INFINITY;
        for (int t = 0; t < T; ++t) {
            const double v = fmax(xn + cnorm[t] - 2.0 * dots[i * T + t], 0.0);
            d2[(long long)t * n + i] = v < cl ? v : cl;
        }
    }
}
}  // namespace sq

extern "C" int sq_kpp_update(const double* dots, const double* xnorm, const double* cnorm, const double* closest,
                             long long n, int T, double* d2, void* stream_) {
    cudaStream_t stream = (cudaStream_t)stream_;
    SQ_REQUIRE(dots && xnorm && cnorm && d2 && n >= 1 && T >= 1, "sq_kpp_update: bad arguments");
    long long blocks = (n + 255) / 256;
    int cap = device_sm_count() * 16;
    if (blocks > cap) blocks = cap;
    kpp_update_kernel<<<(int)blocks, 256, 0, stream>>>(dots, xnorm, cnorm, closest, n, T, d2);
    SQ_LAUNCH_CHECK("kpp_update_kernel");
    return SQ_OK;
}
