// K2 — q-means E-step: fp64 tensor-core distance GEMM fused with the noise
// model and the arg-min, so the n x k distance matrix never reaches HBM.
//
// Replaces labels_estimation (sklearn/cluster/_dmeans.py:732-777), ipe
// (sklearn/QuantumUtility/Utility.py:697-737), amplitude_estimation
// (:442-531) + median_evaluation (:534-572) and select_labels
// (_dmeans.py:2252-2257).  Three modes, as in the reference:
//   EXACT  (delta == 0)               labels = argmin_j ||x_i - c_j||^2, exact
//                                     ties broken uniformly at random (:771-772)
//   WINDOW (delta > 0, !true_distance_estimate)   label drawn uniformly from
//                                     {j : D_ij <= min_j D_ij + delta} (:740-751)
//   IPE    (delta > 0, true_distance_estimate)    D_ij = |x|^2+|c|^2-2*ipe(x,c,delta/2),
//                                     ipe = median of 13 amplitude-estimation
//                                     draws (:753-769), then argmin
// In every mode the reported inertia is sum_i min_j D_ij (:751, :773-775).
// IPE draws the median by one inversion per pair (ipe_lean.cuh) and leaves out, with a certificate from the pair's own
// uniform, the pairs whose estimate certainly exceeds one already sampled in the row: labels, distances and inertia are
// those of sampling every pair.
//
// Layout: X [n, d] and C [k, d] row-major fp64.  One CTA owns BM rows and
// sweeps BN centroids at a time; X and C stream through a STAGES-deep
// cp.async ring in k-chunks of 16, stored as [k/4][row][4] with the row
// XOR-swizzled by (k/4) so both the async writes and the DMMA fragment reads
// are bank-conflict free.  D = X C^T accumulates in registers with
// mma.sync.m8n8k4.f64 (DMMA.8x8x4; tcgen05 has no fp64 kind).
//
// Bit-exact labels against cdist()**2: the GEMM form |c|^2 - 2 x.c rounds
// differently from the reference's direct difference, so rows whose best and
// second-best differ by less than a rounding bound are re-evaluated by
// qmeans_fixup_kernel with the reference's own operation order (sequential
// sum of squared differences, sqrt, square).
#include "common.cuh"
#include "sqb200.h"
#include "fejer.cuh"
#include "ipe.cuh"
#include "ipe_lean.cuh"
#include "binomial.cuh"
#include <cstdlib>
#include <cmath>
#include <cstring>

namespace sq {

enum AssignMode { MODE_EXACT = 0, MODE_WINDOW = 1, MODE_IPE = 2, MODE_COLABS = 3, MODE_STORE = 4 };

struct AssignArgs {
    const double* X;       // [n, d]
    const double* xnorm;   // [n]   ||x_i||^2
    const double* C;       // [k, d]
    const double* cnorm;   // [k]   ||c_j||^2
    const double* rowmin_in;  // [n] optional precomputed min_j(|c_j|^2 - 2 x.c_j) (WINDOW, k > BN)
    int* labels;           // [n]
    double* mindist;       // [n]   min_j D_ij (clamped at 0)
    double* rowmin_out;    // [n]   optional raw min (EXACT)
    int* flag_count;       // [1]   number of near-tie rows
    int* flag_rows;        // [n]   their indices
    double* ipe_scratch;   // [grid][BM*BN] MODE_IPE: each CTA's tile of dot products, then of estimates
    size_t ipe_scratch_bytes;   // (host) capacity of ipe_scratch: bounds the grid
    int ipe_force_replay;       // test hook: every pair takes the draw-by-draw replay path of ipe.cuh
    int ipe_prune;              // 1: pairs that certainly lose the row's arg-min are not sampled (same results)
    int fix_mindist;            // 1: the window fix-up also writes the row's distance (rows the bf16 filter gave up on)
    double* store;         // [n, k] MODE_STORE: x_i.c_j - cnorm[j]
    unsigned long long* colkey;  // [k] MODE_COLABS: max over rows of (bits(|y|) with the sign of y in bit 0)
    long long n;
    int d, k;
    int num_tiles;         // ceil(n / BM)
    int NT, KC;            // centroid tiles, k-chunks
    int mode;
    double delta;
    double tie_tol;        // relative rounding bound for the near-tie test
    unsigned long long seed, offset;
};

template <int BM_, int BN_, int WARPS_M_, int WARPS_N_, int STAGES_, int BK_ = 16, int MIN_CTAS_ = 1>
struct TileCfg {
    static constexpr int BM = BM_, BN = BN_, WARPS_M = WARPS_M_, WARPS_N = WARPS_N_, STAGES = STAGES_;
    static constexpr int BK = BK_, KB = BK / 4, MIN_CTAS = MIN_CTAS_;
    static constexpr int WM = BM / WARPS_M, WN = BN / WARPS_N;
    static constexpr int MI = WM / 8, NI = WN / 8;
    static constexpr int THREADS = 32 * WARPS_M * WARPS_N;
    static constexpr int STAGE_DOUBLES = (BM + BN) * BK;
    static_assert(WM % 8 == 0 && WN % 8 == 0, "warp tile must be a multiple of 8x8");
};

__device__ __forceinline__ double lds_f64(unsigned addr) {
    double v;
    asm volatile("ld.shared.f64 %0, [%1];" : "=d"(v) : "r"(addr));
    return v;
}
__device__ __forceinline__ double dmin(double a, double b) { return b < a ? b : a; }  // no NaN handling needed here

template <class Cfg, bool VEC16, int MODE, bool TMA>
__global__ void __launch_bounds__(Cfg::THREADS, Cfg::MIN_CTAS) qmeans_assign_kernel(AssignArgs a, const __grid_constant__ CUtensorMap tmA,
                                                                          const __grid_constant__ CUtensorMap tmB) {
    constexpr int BM = Cfg::BM, BN = Cfg::BN, MI = Cfg::MI, NI = Cfg::NI, WN_CNT = Cfg::WARPS_N;
    constexpr int STAGES = Cfg::STAGES;
    extern __shared__ __align__(1024) double smem[];           // fragment blocks are 256 B: keep them line-aligned
    double* stages = smem;                                   // STAGES * STAGE_DOUBLES
    double* wmin = stages + STAGES * Cfg::STAGE_DOUBLES;     // [BM][WN_CNT] per-warp row minima
    double* xn_s = wmin + BM * WN_CNT;                       // [BM] ||x||^2 of the tile's rows
    double* thr_s = xn_s + BM;                               // [BM] precomputed row minimum (WINDOW, k > BN)
    double* rowmin_s = thr_s + BM;                           // [BM] this tile's row minimum over all warps
    double* run_m1 = rowmin_s + BM;                          // [BM] running minimum over centroid tiles
    int* candcols = (int*)(run_m1 + BM);                     // [BM][4] first candidate columns of the row, WINDOW
    int* cnt = candcols + 4 * BM;                            // [BM] candidates found in the row
    int* first = cnt + BM;                                   // [BM] lowest candidate column
    int* run_a1 = first + BM;                                // [BM]
    int* run_cnt = run_a1 + BM;                              // [BM]
    double* cn_s = (double*)(run_cnt + BM);                  // [NT*BN] centroid norms (+inf past k)
    unsigned long long* full_bar = (unsigned long long*)(cn_s + a.NT * BN);  // [STAGES] TMA completion barriers
    int* ipe_cnt = (int*)(full_bar + STAGES);                // [2] MODE_IPE: pairs listed for pass 2a / for the general estimator
    unsigned* keep_s = (unsigned*)(ipe_cnt + 2);             // [BM*BN/32] MODE_IPE: bit (row*BN + col) = the pair's value comes back from the workspace

    const int tid = threadIdx.x, lane = tid & 31, warp = tid >> 5;
    const int wm = warp / WN_CNT, wn = warp % WN_CNT;
    const int lq = lane >> 2, lr4 = lane & 3;

    for (int j = tid; j < a.NT * BN; j += Cfg::THREADS) cn_s[j] = (j < a.k) ? a.cnorm[j] : INFINITY;
    if (TMA) {
        if (tid == 0) {
            for (int s = 0; s < STAGES; ++s) mbar_init(&full_bar[s], 1);
            asm volatile("fence.mbarrier_init.release.cluster;" ::: "memory");
            asm volatile("fence.proxy.async.shared::cta;" ::: "memory");
        }
        __syncthreads();
    }

    // tiles owned by this CTA: blockIdx.x, + gridDim.x, ...
    const int my_tiles = (a.num_tiles - (int)blockIdx.x + (int)gridDim.x - 1) / (int)gridDim.x;
    const int per_tile = a.NT * a.KC;
    const int total = my_tiles * per_tile;   // < 2^31: checked on the host

    // ---- loader state: every thread copies the same (row offset, k offset) piece of each pass ----
    constexpr int PPR = VEC16 ? Cfg::BK / 2 : Cfg::BK;  // pieces per row per chunk
    constexpr int RPI = Cfg::THREADS / PPR;        // rows covered by one pass of the CTA
    constexpr int NA = (BM + RPI - 1) / RPI, NB = (BN + RPI - 1) / RPI;
    static_assert(RPI % 4 == 0, "row swizzle needs passes aligned to 4 rows");
    static_assert(MI <= 4, "the multi-candidate path selects accumulator rows by hand for MI <= 4");
    const int r_t = tid / PPR;
    const int kk = VEC16 ? (tid % PPR) * 2 : (tid % PPR);
    const int kb_t = kk >> 2;
    const int dstA0 = ((kb_t * BM + (r_t ^ kb_t)) << 2) + (kk & 3);
    const int dstB0 = BM * Cfg::BK + ((kb_t * BN + (r_t ^ kb_t)) << 2) + (kk & 3);
    const long long rpi_d = (long long)RPI * a.d;
    int p_left = total;
    long long p_tile = blockIdx.x;
    int p_nt = 0, p_kc = 0, p_stage = 0;
    const double* a_ptr = a.X + (p_tile * BM + r_t) * (long long)a.d + kk;
    long long a_rem = a.n - (p_tile * BM + r_t);   // rows left from this thread's first row
    const double* b_ptr = a.C + (long long)r_t * a.d + kk;
    int b_rem = a.k - r_t;

    auto issue = [&]() {
        if (TMA) {
            // one thread hands the whole chunk to the TMA unit: per k-block of 4 one box {4, rows} for X and
            // one for C, landing as dense [k/4][row][4] slabs; rows / columns past the end arrive as zeros
            if (warp == 0 && tid == 0 && p_left > 0) {
                double* st = stages + p_stage * Cfg::STAGE_DOUBLES;
                mbar_expect_tx(&full_bar[p_stage], Cfg::STAGE_DOUBLES * 8);
                const int k0 = p_kc * Cfg::BK;
#pragma unroll
                for (int kb = 0; kb < Cfg::KB; ++kb) {
                    tma_load_2d(st + kb * BM * 4, &tmA, k0 + kb * 4, (int)(p_tile * BM), &full_bar[p_stage]);
                    tma_load_2d(st + BM * Cfg::BK + kb * BN * 4, &tmB, k0 + kb * 4, p_nt * BN, &full_bar[p_stage]);
                }
                --p_left;
                p_stage = (p_stage + 1 == STAGES) ? 0 : p_stage + 1;
                if (++p_kc == a.KC) {
                    p_kc = 0;
                    if (++p_nt == a.NT) { p_nt = 0; p_tile += gridDim.x; }
                }
            }
            return;
        }
        if (p_left > 0) {
            double* st = stages + p_stage * Cfg::STAGE_DOUBLES;
            const int koff = p_kc * Cfg::BK;
            const bool kok = koff + kk < a.d;
#pragma unroll
            for (int i = 0; i < NA; ++i) {
                if (BM % RPI == 0 || r_t + i * RPI < BM) {
                    bool ok = kok && (long long)(i * RPI) < a_rem;
                    const double* src = ok ? a_ptr + i * rpi_d + koff : a.X;
                    if (VEC16) cp_async16(st + dstA0 + i * RPI * 4, src, ok); else cp_async8(st + dstA0 + i * RPI * 4, src, ok);
                }
            }
#pragma unroll
            for (int i = 0; i < NB; ++i) {
                if (BN % RPI == 0 || r_t + i * RPI < BN) {
                    bool ok = kok && i * RPI < b_rem;
                    const double* src = ok ? b_ptr + i * rpi_d + koff : a.C;
                    if (VEC16) cp_async16(st + dstB0 + i * RPI * 4, src, ok); else cp_async8(st + dstB0 + i * RPI * 4, src, ok);
                }
            }
            --p_left;
            p_stage = (p_stage + 1 == STAGES) ? 0 : p_stage + 1;
            if (++p_kc == a.KC) {
                p_kc = 0;
                if (++p_nt == a.NT) {
                    p_nt = 0;
                    p_tile += gridDim.x;
                    a_ptr = a.X + (p_tile * BM + r_t) * (long long)a.d + kk;
                    a_rem = a.n - (p_tile * BM + r_t);
                }
                b_ptr = a.C + ((long long)p_nt * BN + r_t) * a.d + kk;
                b_rem = a.k - (p_nt * BN + r_t);
            }
        }
        cp_async_commit();
    };

    for (int s = 0; s < STAGES - 1; ++s) issue();

    double acc[MI][NI][2];
#pragma unroll
    for (int i = 0; i < MI; ++i)
#pragma unroll
        for (int j = 0; j < NI; ++j) acc[i][j][0] = acc[i][j][1] = 0.0;

    // MODE_COLABS keeps, per owned column, the largest |x.v - mu.v| seen over all of this CTA's tiles
    unsigned long long colbest[NI][2];
#pragma unroll
    for (int j = 0; j < NI; ++j) colbest[j][0] = colbest[j][1] = 0ull;

    // fragment offsets inside a stage (doubles); kb only permutes rows inside aligned groups of 4
    const int fragA = ((wm * Cfg::WM + lq) << 2) + lr4;
    const int fragB = BM * Cfg::BK + ((wn * Cfg::WN + lq) << 2) + lr4;

    int kc = 0, nt = 0, tl = 0, c_stage = 0;
    unsigned c_phase = 0;
    const unsigned stages_u32 = (unsigned)__cvta_generic_to_shared(stages);
    for (int q = 0; q < total; ++q) {
        if (TMA) {
            // The stage about to be refilled was read in the previous iteration.  Only the warp that issues
            // the TMA has to wait for everybody to be done with it: the others just say so and go on
            // (one named barrier per stage of the ring: a warp can run at most STAGES - 1 iterations ahead of the
            // slowest one - it needs the load issued after that barrier completed - so an arrival for iteration
            // q + STAGES can never be counted in iteration q).
            if (warp == 0) {
                asm volatile("barrier.sync %0, %1;" ::"r"(1 + c_stage), "r"(Cfg::THREADS) : "memory");
                issue();
            } else {
                asm volatile("barrier.arrive %0, %1;" ::"r"(1 + c_stage), "r"(Cfg::THREADS) : "memory");
            }
            mbar_wait(&full_bar[c_stage], c_phase);
        } else {
            cp_async_wait<STAGES - 2>();
            __syncthreads();
            issue();
        }

        const unsigned st_u32 = stages_u32 + (unsigned)c_stage * (Cfg::STAGE_DOUBLES * 8);
        const unsigned fa_u32 = st_u32 + fragA * 8, fb_u32 = st_u32 + fragB * 8;
        if (++c_stage == STAGES) { c_stage = 0; c_phase ^= 1; }
#pragma unroll
        for (int kb = 0; kb < Cfg::KB; ++kb) {
            double af[MI], bf[NI];
            // (row ^ kb) with row = base + 8*i + lq: only lq's low two bits change
            const int swz = TMA ? 0 : ((lq ^ kb) - lq) << 5;   // bytes; TMA slabs are dense, cp.async slabs XOR-swizzled
#pragma unroll
            for (int i = 0; i < MI; ++i) af[i] = lds_f64(fa_u32 + swz + ((kb * BM + i * 8) << 5));
#pragma unroll
            for (int j = 0; j < NI; ++j) bf[j] = lds_f64(fb_u32 + swz + ((kb * BN + j * 8) << 5));
#pragma unroll
            for (int i = 0; i < MI; ++i)
#pragma unroll
                for (int j = 0; j < NI; ++j) dmma884(acc[i][j][0], acc[i][j][1], af[i], bf[j]);
        }

        if (kc == a.KC - 1 && MODE == MODE_COLABS) {
            // column-wise |max| of Y = X V^T - mu V^T (svd_flip's arg-max over all rows); NT must be 1
            const long long tile = (long long)blockIdx.x + (long long)tl * gridDim.x;
            const long long r0 = tile * BM;
            const int colbase = wn * Cfg::WN + lr4 * 2;
#pragma unroll
            for (int j = 0; j < NI; ++j) {
                const double2 mv = *reinterpret_cast<const double2*>(cn_s + colbase + j * 8);
#pragma unroll
                for (int i = 0; i < MI; ++i) {
                    if (r0 + wm * Cfg::WM + i * 8 + lq < a.n) {
                        double y0 = acc[i][j][0] - mv.x, y1 = acc[i][j][1] - mv.y;
                        unsigned long long k0 = ((unsigned long long)__double_as_longlong(fabs(y0)) & ~1ull) | (y0 < 0.0 ? 1ull : 0ull);
                        unsigned long long k1 = ((unsigned long long)__double_as_longlong(fabs(y1)) & ~1ull) | (y1 < 0.0 ? 1ull : 0ull);
                        colbest[j][0] = k0 > colbest[j][0] ? k0 : colbest[j][0];
                        colbest[j][1] = k1 > colbest[j][1] ? k1 : colbest[j][1];
                    }
                    acc[i][j][0] = acc[i][j][1] = 0.0;
                }
            }
        } else if (kc == a.KC - 1 && MODE == MODE_STORE) {
            // plain projection: Y[i, j] = x_i.c_j - offset[j] written out (qPCA.transform, _base.py:117-128)
            const long long tile = (long long)blockIdx.x + (long long)tl * gridDim.x;
            const long long r0 = tile * BM;
            const int colbase = nt * BN + wn * Cfg::WN + lr4 * 2;
#pragma unroll
            for (int i = 0; i < MI; ++i) {
                const long long gr = r0 + wm * Cfg::WM + i * 8 + lq;
#pragma unroll
                for (int j = 0; j < NI; ++j) {
                    const int col = colbase + j * 8;
                    if (gr < a.n) {
                        if (col < a.k) a.store[gr * a.k + col] = acc[i][j][0] - cn_s[col];
                        if (col + 1 < a.k) a.store[gr * a.k + col + 1] = acc[i][j][1] - cn_s[col + 1];
                    }
                    acc[i][j][0] = acc[i][j][1] = 0.0;
                }
            }
        } else if (kc == a.KC - 1) {
            // ------------------------------ epilogue ------------------------------
            // v = |c|^2 - 2 x.c per element; the row minimum is found first, then the set of
            // "candidates" v <= min + thr: thr = delta (WINDOW), a rounding bound (EXACT: more than
            // one candidate sends the row to the exact fix-up) or 0 (IPE).  Rows with a single
            // candidate - the common case - need no random number at all.
            const long long tile = (long long)blockIdx.x + (long long)tl * gridDim.x;
            const long long r0 = tile * BM;
            const int n0 = nt * BN;
            const bool last_nt = (nt == a.NT - 1);
            const int colbase = n0 + wn * Cfg::WN + lr4 * 2;

            if (tid < BM) {
                long long gr = r0 + tid;
                xn_s[tid] = (gr < a.n) ? a.xnorm[gr] : 0.0;
                if (MODE == MODE_WINDOW && a.rowmin_in) thr_s[tid] = (gr < a.n) ? a.rowmin_in[gr] : 0.0;
                if (MODE != MODE_WINDOW || nt == 0) { cnt[tid] = 0; first[tid] = 0x7fffffff; }
            }
            if (MODE == MODE_IPE) {
                // The accumulator (x.c) becomes (noisy squared distance - |x|^2), or +inf for a pair that certainly
                // loses the row's arg-min.  The tile of dot products is parked in this CTA's slice of the workspace
                // (L2-resident), so that the sampler runs with the whole register file instead of beside 2*MI*NI
                // live accumulators, and the pairs are dealt to the lanes as a flat list (no idle lanes for k < BN).
                double* sc = a.ipe_scratch + (size_t)blockIdx.x * (BM * BN + BM * BN / 2);   // dots, then the pass-2 list
                const int cl0 = wn * Cfg::WN + lr4 * 2;
#pragma unroll
                for (int i = 0; i < MI; ++i)
#pragma unroll
                    for (int j = 0; j < NI; ++j)
                        *reinterpret_cast<double2*>(sc + (wm * Cfg::WM + i * 8 + lq) * BN + cl0 + j * 8) =
                            make_double2(acc[i][j][0], acc[i][j][1]);
                for (int w = tid; w < BM * BN / 32; w += Cfg::THREADS) keep_s[w] = 0u;
                const int kk = min(BN, a.k - n0);                  // live columns of this centroid tile
                const int rows = (int)min((long long)BM, a.n - r0);
                int* list = reinterpret_cast<int*>(sc + BM * BN);
                if (tid == 0) { ipe_cnt[0] = 0; ipe_cnt[1] = 0; }
                const double eps = 0.5 * a.delta;
                const bool prune = a.ipe_prune && !a.ipe_force_replay;
                double* estar_s = thr_s;                           // [BM] (WINDOW's array, free here)
                int* cand_s = candcols;                            // [BM] (WINDOW's array, free here)
                if (prune) {
                    // Pass 0: the row's most promising centroid of this tile (smallest true distance) is sampled first
                    // and in full; its estimate - or a smaller one from an earlier centroid tile - is what every other
                    // pair of the row has to beat (ipe_out_possible).
                    // (any column would do as the reference, so the arg-min is taken on keys: |c|^2 - 2 x.c with the
                    // column in its low 8 mantissa bits, straight from the accumulators)
                    static_assert(BN <= 256, "the candidate key carries the column in 8 bits");
#pragma unroll
                    for (int i = 0; i < MI; ++i) {
                        double key = INFINITY;
#pragma unroll
                        for (int j = 0; j < NI; ++j)
#pragma unroll
                            for (int e = 0; e < 2; ++e) {
                                const int col = cl0 + j * 8 + e;
                                const double v = fma(-2.0, acc[i][j][e], cn_s[n0 + col]);          // +inf past k
                                const double kv = __longlong_as_double((__double_as_longlong(v) & ~0xffLL) | (long long)col);
                                key = dmin(key, (v < INFINITY) ? kv : INFINITY);
                            }
                        key = dmin(key, __shfl_xor_sync(0xffffffffu, key, 1));
                        key = dmin(key, __shfl_xor_sync(0xffffffffu, key, 2));
                        if (lr4 == 0) wmin[(wm * Cfg::WM + i * 8 + lq) * WN_CNT + wn] = key;
                    }
                    __syncthreads();
                    {
                        constexpr int ENTRIES = BM * WN_CNT;
                        for (int e0 = 0; e0 < ENTRIES; e0 += Cfg::THREADS) {
                            const int e = e0 + tid;
                            double m = (e < ENTRIES) ? wmin[e] : INFINITY;
#pragma unroll
                            for (int o = 1; o < WN_CNT; o <<= 1) m = dmin(m, __shfl_xor_sync(0xffffffffu, m, o));
                            if (e < ENTRIES && (e % WN_CNT) == 0)
                                cand_s[e / WN_CNT] = (m < INFINITY) ? (int)(__double_as_longlong(m) & 0xffLL) : 0;
                        }
                    }
                    __syncthreads();
                    static_assert(BM <= Cfg::THREADS, "one thread per row");
                    if (tid < rows) {                              // full warps: the chain is as long, the issue slots are not
                        const int row = tid, cl = cand_s[row];
                        const double xn = xn_s[row];
                        IpeGrid g;
                        ipe_grid(sc[row * BN + cl], xn, cn_s[n0 + cl], eps, g);
                        double e = INFINITY;
                        if (g.kind == 2) {
                            const double V = ipe_pair_uniform(a.seed, a.offset + (unsigned long long)(r0 + row), (unsigned)(n0 + cl));
                            double med = 0.0;
                            int st = ipe_lean(g, V, med);
                            if (st == IPE_TIGHTEN) st = ipe_tighten(g, V, med);
                            if (st == IPE_BRUTE) { med = ipe_brute(g, V); st = IPE_DONE; }
                            if (st == IPE_DONE) e = ipe_minus(ipe_value(g, med), xn);
                            else list[BM * BN - 1 - atomicAdd(ipe_cnt + 1, 1)] = row * BN + cl;       // the general estimator's
                        } else {
                            e = ipe_minus((g.kind == 1) ? ipe_value(g, fmin(g.j0, g.M - g.j0)) : 0.0, xn);
                        }
                        if (e < INFINITY) sc[row * BN + cl] = e;
                        atomicOr(&keep_s[(row * BN + cl) >> 5], 1u << ((row * BN + cl) & 31));
                        estar_s[row] = fmin(e, nt > 0 ? run_m1[row] : INFINITY);
                    }
                }
                __syncthreads();  // list counter, estar_s, cand_s
                // Pass 1, one pair per lane, no divergence: grid size and the pair's uniform; a pair whose estimate is
                // certainly above the row's reference estimate is out (+inf), the others are listed (front of the
                // list) for pass 2a.  With the replay hook everything goes to the general estimator (back of the list).
                // (the dot product of the next round is fetched while this round's test is computed)
                constexpr int LIST_CAP = BM * BN;
                // (one unit = the eight columns 8c .. 8c + 7 of a row: they share one block of the generator, whose 16
                // bits per pair bound the pair's uniform from below - enough for the certificate; the dot products of
                // the next round are fetched while this round's tests are computed)
                static_assert(BN % 8 == 0, "units of eight columns");
                const int hk = (kk + 7) >> 3, units = rows * hk;
                const int urounds = (units + Cfg::THREADS - 1) / Cfg::THREADS;
                int row = 0, c8 = 0;
                bool ok = tid < units;
                if (ok) { row = tid / hk; c8 = tid - row * hk; }
                const int drow = Cfg::THREADS / hk, dc8 = Cfg::THREADS - drow * hk;     // one round further on
                double2 dots[4];
#pragma unroll
                for (int h = 0; h < 4; ++h) dots[h] = *reinterpret_cast<const double2*>(sc + row * BN + 8 * c8 + 2 * h);
#pragma unroll 1
                for (int it = 0; it < urounds; ++it) {
                    const bool ok2 = (it + 1) * Cfg::THREADS + tid < units;
                    int row2 = row + drow, c82 = c8 + dc8;
                    if (c82 >= hk) { c82 -= hk; ++row2; }
                    if (!ok2) { row2 = 0; c82 = 0; }
                    double2 dots2[4];
#pragma unroll
                    for (int h = 0; h < 4; ++h) dots2[h] = *reinterpret_cast<const double2*>(sc + row2 * BN + 8 * c82 + 2 * h);
                    const int cl = 8 * c8;
                    const double xn = xn_s[row];
                    const double es = prune ? estar_s[row] : INFINITY;
                    const int cand = prune ? cand_s[row] : -1;
                    uint32_t hb[4] = {0u, 0u, 0u, 0u};
                    if (prune && ok) ipe_oct_bits(a.seed, a.offset + (unsigned long long)(r0 + row), (unsigned)(n0 + cl), hb);
                    unsigned want = 0, keep = 0;                   // want: bit e = column cl + e is to be sampled in pass 2a
#pragma unroll
                    for (int e = 0; e < 8; ++e) {
                        const double dot = (e & 1) ? dots[e >> 1].y : dots[e >> 1].x;
                        if (ok && cl + e < kk && cl + e != cand) {
                            const double cn = cn_s[n0 + cl + e];
                            if (!(xn + cn > 0.0)) {
                                sc[row * BN + cl + e] = ipe_minus(0.0, xn);                // |x|^2 + |c|^2 = 0: the estimate is 0
                                keep |= 1u << e;
                            } else if (a.ipe_force_replay) {
                                list[LIST_CAP - 1 - atomicAdd(ipe_cnt + 1, 1)] = row * BN + cl + e;
                                keep |= 1u << e;
                            } else {
                                double lhs, rhs;
                                const bool out = prune && ipe_out_possible(dot, xn, cn, eps, es, lhs, rhs) &&
                                                 (double)ipe_oct_high(hb, (unsigned)e) * (1.0 / 65536.0) * lhs > rhs;
                                if (!out) want |= 1u << e;
                            }
                        }
                    }
                    // the listed pairs: one counter bump per warp, one bit-mask update per unit
                    {
                        const int mine = __popc(want);
                        int incl = mine;
#pragma unroll
                        for (int o = 1; o < 32; o <<= 1) {
                            const int y = __shfl_up_sync(0xffffffffu, incl, o);
                            if (lane >= o) incl += y;
                        }
                        const int total = __shfl_sync(0xffffffffu, incl, 31);
                        if (total) {
                            int base = 0;
                            if (lane == 31) base = atomicAdd(ipe_cnt, total);
                            base = __shfl_sync(0xffffffffu, base, 31) + incl - mine;
                            unsigned w = want;
                            while (w) {
                                const int e = __ffs(w) - 1;
                                w &= w - 1;
                                list[base++] = row * BN + cl + e;
                            }
                        }
                        keep |= want;
                        if (keep) atomicOr(&keep_s[(row * BN + cl) >> 5], keep << ((row * BN + cl) & 31));
                    }
                    row = row2; c8 = c82; ok = ok2;
#pragma unroll
                    for (int h = 0; h < 4; ++h) dots[h] = dots2[h];
                }
                __syncthreads();
                const int nres = ipe_cnt[0];
                // Pass 2a: the median of the 13 amplitude-estimation outcomes by one inversion (ipe_lean.cuh), one
                // listed pair per lane, packed into as few warps as possible (a short list is one dependent chain
                // either way; full warps leave the issue slots to the CTA next door).  A pair the inversion declines
                // before looking at its uniform (window touching M/2 on a large grid) moves to the back of the list for
                // the general estimator.
#pragma unroll 1
                for (int li = tid; li < nres; li += Cfg::THREADS) {
                    const int pos = list[li], rowd = pos / BN, cld = pos - rowd * BN;
                    const double xn = xn_s[rowd];
                    IpeGrid g;
                    ipe_grid(sc[pos], xn, cn_s[n0 + cld], eps, g);
                    double v;
                    if (g.kind == 1) {
                        v = ipe_value(g, fmin(g.j0, g.M - g.j0));      // weight 1 at distance 0 (Utility.py:500-501)
                    } else {
                        const double V = ipe_pair_uniform(a.seed, a.offset + (unsigned long long)(r0 + rowd), (unsigned)(n0 + cld));
                        double med = 0.0;
                        int st = ipe_lean(g, V, med);
                        if (st == IPE_TIGHTEN) st = ipe_tighten(g, V, med);
                        if (st == IPE_BRUTE) { med = ipe_brute(g, V); st = IPE_DONE; }
                        if (st != IPE_DONE) {
                            list[LIST_CAP - 1 - atomicAdd(ipe_cnt + 1, 1)] = pos;
                            continue;
                        }
                        v = ipe_value(g, med);
                    }
                    sc[pos] = ipe_minus(v, xn);
                }
                __syncthreads();
                const int ngen = ipe_cnt[1];
                // Pass 2b: the general estimator (warp-cooperative: every lane of a warp calls it together, and its time
                // grows with the number of live lanes - so the entries are dealt to the warps first here too)
#pragma unroll 1
                for (int it = 0; it * Cfg::THREADS < ngen; ++it) {
                    const int li = it * Cfg::THREADS + lane * (Cfg::THREADS / 32) + warp;
                    const bool okd = li < ngen;
                    const int pos = okd ? list[LIST_CAP - 1 - li] : 0;
                    const int rowd = pos / BN, cld = pos - rowd * BN;
                    const double xn = xn_s[rowd];
                    bool defer;
                    const double v = ipe_estimate_warp<false>(sc[pos], xn, okd ? cn_s[n0 + cld] : 0.0, eps, okd,
                                                              a.seed, a.offset + (unsigned long long)(r0 + rowd),
                                                              (unsigned)(n0 + cld), a.ipe_force_replay != 0, defer);
                    if (okd) sc[pos] = ipe_minus(v, xn);
                }
                if (ngen > 0) __syncthreads();                     // (the same for every thread; usually there is nothing to wait for)
#pragma unroll
                for (int i = 0; i < MI; ++i) {
                    const int row = wm * Cfg::WM + i * 8 + lq;
#pragma unroll
                    for (int j = 0; j < NI; ++j) {
                        // (only the pairs that were sampled come back; the others are out: +inf)
                        const int lc = cl0 + j * 8;
                        const unsigned kb2 = (keep_s[(row * BN + lc) >> 5] >> ((row * BN + lc) & 31)) & 3u;
                        acc[i][j][0] = (kb2 & 1u) ? sc[row * BN + lc] : INFINITY;
                        acc[i][j][1] = (kb2 & 2u) ? sc[row * BN + lc + 1] : INFINITY;
                    }
                }
            }
            // ---- phase A: v in place, per-warp row minima ----
#pragma unroll
            for (int j = 0; j < NI; ++j) {
                const double2 cn2 = *reinterpret_cast<const double2*>(cn_s + colbase + j * 8);
#pragma unroll
                for (int i = 0; i < MI; ++i) {
                    if (MODE != MODE_IPE) {   // (IPE already holds noisy distance - |x|^2)
                        acc[i][j][0] = fma(-2.0, acc[i][j][0], cn2.x);
                        acc[i][j][1] = fma(-2.0, acc[i][j][1], cn2.y);
                    }
                }
            }
#pragma unroll
            for (int i = 0; i < MI; ++i) {
                double m = dmin(acc[i][0][0], acc[i][0][1]);
#pragma unroll
                for (int j = 1; j < NI; ++j) m = dmin(m, dmin(acc[i][j][0], acc[i][j][1]));
                m = dmin(m, __shfl_xor_sync(0xffffffffu, m, 1));
                m = dmin(m, __shfl_xor_sync(0xffffffffu, m, 2));
                if (lr4 == 0) wmin[(wm * Cfg::WM + i * 8 + lq) * WN_CNT + wn] = m;
            }
            __syncthreads();
            // ---- phase B: row minimum over the warps (one entry per thread, segmented shuffle), candidates ----
            {
                constexpr int ENTRIES = BM * WN_CNT;
                static_assert((WN_CNT & (WN_CNT - 1)) == 0 && WN_CNT <= 32, "warps along N must be a power of two");
                for (int e0 = 0; e0 < ENTRIES; e0 += Cfg::THREADS) {
                    int e = e0 + tid;
                    double m = (e < ENTRIES) ? wmin[e] : INFINITY;
#pragma unroll
                    for (int o = 1; o < WN_CNT; o <<= 1) m = dmin(m, __shfl_xor_sync(0xffffffffu, m, o));
                    if (e < ENTRIES && (e % WN_CNT) == 0) rowmin_s[e / WN_CNT] = m;
                }
            }
            __syncthreads();
#pragma unroll
            for (int i = 0; i < MI; ++i) {
                const int row = wm * Cfg::WM + i * 8 + lq;
                const double tmin = rowmin_s[row];
                double thr;
                if (MODE == MODE_WINDOW) thr = (a.rowmin_in ? thr_s[row] : tmin) + a.delta;
                else if (MODE == MODE_EXACT) thr = tmin + fmax(a.tie_tol, 0.0) * (xn_s[row] + 2.0 * fabs(tmin));
                else thr = tmin;
                unsigned mask = 0;
#pragma unroll
                for (int j = 0; j < NI; ++j)
#pragma unroll
                    for (int e = 0; e < 2; ++e) mask |= (acc[i][j][e] <= thr ? 1u : 0u) << (j * 2 + e);
                if (mask) {
                    if (MODE == MODE_WINDOW) {
                        // remember up to four candidate columns per row; rows with more go to the fix-up list
                        while (mask) {
                            int b0 = __ffs(mask) - 1;
                            mask &= mask - 1;
                            int slot = atomicAdd(&cnt[row], 1);
                            if (slot < 4) candcols[row * 4 + slot] = colbase + (b0 >> 1) * 8 + (b0 & 1);
                        }
                    } else {
                        int b0 = __ffs(mask) - 1;
                        atomicAdd(&cnt[row], __popc(mask));
                        atomicMin(&first[row], colbase + (b0 >> 1) * 8 + (b0 & 1));
                    }
                }
            }
            __syncthreads();
            // ---- phase C: one thread per row folds the tile into the running state and writes out ----
            if (tid < BM) {
                const int row = tid;
                const long long gr = r0 + row;
                const double tmin = rowmin_s[row];
                const double xn = xn_s[row];
                if (MODE == MODE_WINDOW) {
                    if (last_nt && gr < a.n) {
                        // uniform pick among the candidates (reference: random.choice, _dmeans.py:2252-2257)
                        const int c = cnt[row];
                        int c0 = candcols[row * 4], c1 = candcols[row * 4 + 1], c2 = candcols[row * 4 + 2],
                            c3 = candcols[row * 4 + 3];
                        int label = c0;
                        if (c > 4) {
                            int slot = atomicAdd(a.flag_count, 1);
                            a.flag_rows[slot] = (int)gr;          // redone by qmeans_window_fixup_kernel
                        } else if (c > 1) {
                            // arrival order in the slots is arbitrary: sort, then take the r-th
                            if (c < 4) c3 = 0x7fffffff;
                            if (c < 3) c2 = 0x7fffffff;
                            int t;
                            if (c0 > c1) { t = c0; c0 = c1; c1 = t; }
                            if (c2 > c3) { t = c2; c2 = c3; c3 = t; }
                            if (c0 > c2) { t = c0; c0 = c2; c2 = t; }
                            if (c1 > c3) { t = c1; c1 = c3; c3 = t; }
                            if (c1 > c2) { t = c1; c1 = c2; c2 = t; }
                            Philox rng(a.seed, a.offset + (unsigned long long)gr, 0x51CEu, 0u);
                            int r = (int)(rng.uniform() * c);
                            if (r >= c) r = c - 1;
                            label = r == 0 ? c0 : r == 1 ? c1 : r == 2 ? c2 : c3;
                        }
                        a.labels[gr] = label;
                        const double base = a.rowmin_in ? thr_s[row] : tmin;
                        a.mindist[gr] = fmax(xn + base, 0.0);
                    }
                } else {
                    const int c = cnt[row], f = first[row];
                    double m1 = tmin;
                    int a1 = f, rc = c;
                    if (nt > 0) {
                        const double pm = run_m1[row];
                        const double tol = (MODE == MODE_EXACT) ? fmax(a.tie_tol, 0.0) * (xn + 2.0 * fabs(fmin(pm, tmin))) : 0.0;
                        if (tmin < pm - tol) { /* new tile wins outright */ }
                        else if (tmin <= pm + tol) {        // indistinguishable: count both (over-counting only flags more rows)
                            rc = c + run_cnt[row];
                            if (!(tmin < pm || (tmin == pm && f < run_a1[row]))) { m1 = pm; a1 = run_a1[row]; }
                        } else { m1 = pm; a1 = run_a1[row]; rc = run_cnt[row]; }
                    }
                    run_m1[row] = m1; run_a1[row] = a1; run_cnt[row] = rc;
                    if (last_nt && gr < a.n) {
                        a.labels[gr] = a1;
                        a.mindist[gr] = fmax(xn + m1, 0.0);
                        if (a.rowmin_out) a.rowmin_out[gr] = m1;
                        if (MODE == MODE_EXACT && rc > 1 && a.tie_tol >= 0.0) {
                            int slot = atomicAdd(a.flag_count, 1);
                            a.flag_rows[slot] = (int)gr;  // n < 2^31 per call (checked on the host)
                        }
                    }
                }
            }
#pragma unroll
            for (int i = 0; i < MI; ++i)
#pragma unroll
                for (int j = 0; j < NI; ++j) acc[i][j][0] = acc[i][j][1] = 0.0;
        }

        if (++kc == a.KC) { kc = 0; if (++nt == a.NT) { nt = 0; ++tl; } }
    }
    cp_async_wait<0>();
    if (MODE == MODE_COLABS) {
#pragma unroll
        for (int j = 0; j < NI; ++j)
#pragma unroll
            for (int e = 0; e < 2; ++e) {
                unsigned long long v = colbest[j][e];
#pragma unroll
                for (int o = 4; o < 32; o <<= 1) {
                    unsigned long long w = __shfl_xor_sync(0xffffffffu, v, o);
                    v = w > v ? w : v;
                }
                int col = wn * Cfg::WN + lr4 * 2 + j * 8 + e;
                if (lq == 0 && col < a.k) atomicMax(&a.colkey[col], v);
            }
    }
}

// ---- near-tie rows: the reference's own arithmetic ---------------------------
// cdist(X, C, "euclidean")**2 (_dmeans.py:736-737): s = sum_t (x_t - c_t)^2 in
// index order without fused multiply-add, then sqrt, then square.  One warp
// per flagged row, lanes over centroids; exact ties broken uniformly (:772).
__global__ void qmeans_fixup_kernel(AssignArgs a) {
    const int nflag = *a.flag_count;
    const int lane = threadIdx.x & 31;
    const int wglobal = (blockIdx.x * blockDim.x + threadIdx.x) >> 5;
    const int nwarps = (gridDim.x * blockDim.x) >> 5;
    for (int f = wglobal; f < nflag; f += nwarps) {
        const long long gr = a.flag_rows[f];
        const double* x = a.X + gr * (long long)a.d;
        double best = INFINITY;
        for (int j = lane; j < a.k; j += 32) {
            const double* c = a.C + (long long)j * a.d;
            double s = 0.0;
            for (int t = 0; t < a.d; ++t) {
                double df = __dsub_rn(x[t], c[t]);
                s = __dadd_rn(s, __dmul_rn(df, df));
            }
            double r = __dsqrt_rn(s);
            double dd = __dmul_rn(r, r);
            best = fmin(best, dd);
        }
#pragma unroll
        for (int o = 16; o > 0; o >>= 1) best = fmin(best, __shfl_xor_sync(0xffffffffu, best, o));
        // count exact ties and pick the r-th of them uniformly
        int ties = 0;
        for (int j0 = 0; j0 < a.k; j0 += 32) {
            int j = j0 + lane;
            bool hit = false;
            if (j < a.k) {
                const double* c = a.C + (long long)j * a.d;
                double s = 0.0;
                for (int t = 0; t < a.d; ++t) {
                    double df = __dsub_rn(x[t], c[t]);
                    s = __dadd_rn(s, __dmul_rn(df, df));
                }
                double r = __dsqrt_rn(s);
                hit = (__dmul_rn(r, r) == best);
            }
            ties += __popc(__ballot_sync(0xffffffffu, hit));
        }
        Philox rng(a.seed, a.offset + (unsigned long long)gr, 0x71E5u, 0u);
        int pick = (int)(rng.uniform() * ties);
        if (pick >= ties) pick = ties - 1;
        int label = -1, seen = 0;
        for (int j0 = 0; j0 < a.k && label < 0; j0 += 32) {
            int j = j0 + lane;
            bool hit = false;
            if (j < a.k) {
                const double* c = a.C + (long long)j * a.d;
                double s = 0.0;
                for (int t = 0; t < a.d; ++t) {
                    double df = __dsub_rn(x[t], c[t]);
                    s = __dadd_rn(s, __dmul_rn(df, df));
                }
                double r = __dsqrt_rn(s);
                hit = (__dmul_rn(r, r) == best);
            }
            unsigned m = __ballot_sync(0xffffffffu, hit);
            int cnt = __popc(m);
            if (pick < seen + cnt) {
                int want = pick - seen;
                for (int b = 0; b < 32; ++b)
                    if (m >> b & 1) { if (want == 0) { label = j0 + b; break; } --want; }
            }
            seen += cnt;
        }
        if (lane == 0) {
            a.labels[gr] = label;
            a.mindist[gr] = best;
        }
    }
}

// ---- delta-window rows with more than four candidates ------------------------------
// Rare (wide windows): one warp recomputes the row's |c|^2 - 2 x.c, finds the candidates
// {v <= min + delta} and picks one uniformly (labels_estimation's delta_means1, _dmeans.py:742-748).
__device__ __forceinline__ double fixup_v(const double* x, const double* c, double cn, int d) {
    double s = 0.0;
    for (int t = 0; t < d; ++t) s = fma(x[t], c[t], s);
    return cn - 2.0 * s;
}
__global__ void qmeans_window_fixup_kernel(AssignArgs a) {
    const int nflag = *a.flag_count;
    const int lane = threadIdx.x & 31;
    const int wglobal = (blockIdx.x * blockDim.x + threadIdx.x) >> 5;
    const int nwarps = (gridDim.x * blockDim.x) >> 5;
    for (int f = wglobal; f < nflag; f += nwarps) {
        const long long gr = a.flag_rows[f];
        const double* x = a.X + gr * (long long)a.d;
        double vmin = INFINITY;
        int amin = 0;
        for (int j = lane; j < a.k; j += 32) {
            double v = fixup_v(x, a.C + (long long)j * a.d, a.cnorm[j], a.d);
            if (v < vmin) { vmin = v; amin = j; }
        }
#pragma unroll
        for (int o = 16; o > 0; o >>= 1) {
            double ov = __shfl_xor_sync(0xffffffffu, vmin, o);
            int oa = __shfl_xor_sync(0xffffffffu, amin, o);
            if (ov < vmin || (ov == vmin && oa < amin)) { vmin = ov; amin = oa; }
        }
        const double base = a.rowmin_in ? a.rowmin_in[gr] : vmin;
        const double thr = base + a.delta;
        int total = 0;
        for (int j0 = 0; j0 < a.k; j0 += 32) {
            int j = j0 + lane;
            bool hit = j < a.k && fixup_v(x, a.C + (long long)j * a.d, a.cnorm[j], a.d) <= thr;
            total += __popc(__ballot_sync(0xffffffffu, hit));
        }
        int label = amin;
        if (total > 0) {
            // (rows the bf16 filter handed over with four candidates or fewer draw what the main kernel would have drawn)
            Philox rng(a.seed, a.offset + (unsigned long long)gr, 0x51CEu, total > 4 ? 1u : 0u);
            int pick = (int)(rng.uniform() * total);
            if (pick >= total) pick = total - 1;
            int seen = 0;
            bool found = false;
            for (int j0 = 0; j0 < a.k && !found; j0 += 32) {
                int j = j0 + lane;
                bool hit = j < a.k && fixup_v(x, a.C + (long long)j * a.d, a.cnorm[j], a.d) <= thr;
                unsigned m = __ballot_sync(0xffffffffu, hit);
                int c = __popc(m);
                if (pick < seen + c) {
                    int want = pick - seen;
                    for (int b = 0; b < 32; ++b)
                        if (m >> b & 1) { if (want == 0) { label = j0 + b; break; } --want; }
                    found = true;
                }
                seen += c;
            }
        }
        if (lane == 0) {
            a.labels[gr] = label;
            if (a.fix_mindist) a.mindist[gr] = fmax(a.xnorm[gr] + base, 0.0);
        }
    }
}

// ---- ||row||^2 of a row-major matrix (row_norms(X, squared=True), extmath.py:49-78)
__global__ void row_sqnorm_kernel(const double* __restrict__ X, long long n, int d, double* __restrict__ out,
                                  int* __restrict__ clear) {
    if (clear && blockIdx.x == 0 && threadIdx.x == 0) *clear = 0;   // (the E-step's near-tie counter rides along)
    const int lane = threadIdx.x & 31;
    long long w = ((long long)blockIdx.x * blockDim.x + threadIdx.x) >> 5;
    const long long nw = ((long long)gridDim.x * blockDim.x) >> 5;
    for (; w < n; w += nw) {
        const double* x = X + w * (long long)d;
        double s = 0.0;
        if ((d & 1) == 0) {
            const double2* x2 = reinterpret_cast<const double2*>(x);
            for (int t = lane; t < d / 2; t += 32) { double2 v = x2[t]; s = fma(v.x, v.x, s); s = fma(v.y, v.y, s); }
        } else {
            for (int t = lane; t < d; t += 32) { double v = x[t]; s = fma(v, v, s); }
        }
        s = warp_sum(s);
        if (lane == 0) out[w] = s;
    }
}

// deterministic sum of an array: fixed grid, fixed per-block order, second pass by one block
__global__ void sum_partial_kernel(const double* __restrict__ v, long long n, double* __restrict__ part) {
    __shared__ double red[32];
    double s = 0.0;
    for (long long i = (long long)blockIdx.x * blockDim.x + threadIdx.x; i < n; i += (long long)gridDim.x * blockDim.x) s += v[i];
    s = warp_sum(s);
    if ((threadIdx.x & 31) == 0) red[threadIdx.x >> 5] = s;
    __syncthreads();
    if (threadIdx.x == 0) {
        double t = 0.0;
        for (int w = 0; w < (int)(blockDim.x >> 5); ++w) t += red[w];
        part[blockIdx.x] = t;
    }
}
__global__ void sum_final_kernel(const double* __restrict__ part, int np, double* __restrict__ out,
                                 const int* __restrict__ copy_src, int* __restrict__ copy_dst) {
    __shared__ double red[32];
    if (copy_dst && threadIdx.x == 0) *copy_dst = *copy_src;      // (the E-step's count of re-evaluated rows rides along)
    double s = 0.0;
    for (int i = threadIdx.x; i < np; i += blockDim.x) s += part[i];
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
#include "qmeans_filter.cuh"
namespace sq {

// ---------------------------------------------------------------------------
// host side
// ---------------------------------------------------------------------------
struct AssignWs {
    int* flag_count;
    int* flag_rows;
    double* cnorm;
    double* rowmin;
    double* partial;
    double* mindist;
};
static const int SUM_BLOCKS = 1024;
// ipe mode parks one tile of dot products and a list of as many ints (<= 96 KB) per resident CTA behind the fixed part;
// sq_qmeans_assign_workspace_bytes asks for enough for 4 CTAs on each of 160 SMs, less still works (smaller grid)
static const size_t IPE_SCRATCH_BYTES = (size_t)640 * 98304;

static size_t assign_ws_layout(long long n, int k, AssignWs* w, char* base) {
    size_t off = 0;
    auto take = [&](size_t bytes) { size_t o = off; off = align_up(off + bytes, 256); return base ? base + o : nullptr; };
    char* p;
    p = take(256);                           if (w) w->flag_count = (int*)p;
    p = take((size_t)n * sizeof(int));       if (w) w->flag_rows = (int*)p;
    p = take((size_t)k * sizeof(double));    if (w) w->cnorm = (double*)p;
    p = take((size_t)n * sizeof(double));    if (w) w->rowmin = (double*)p;
    p = take((size_t)SUM_BLOCKS * sizeof(double)); if (w) w->partial = (double*)p;
    p = take((size_t)n * sizeof(double));    if (w) w->mindist = (double*)p;
    return off;
}

template <class Cfg>
static size_t assign_smem_bytes() {
    return (size_t)Cfg::STAGES * Cfg::STAGE_DOUBLES * 8 + (size_t)Cfg::BM * Cfg::WARPS_N * 8 +
           (size_t)Cfg::BM * (8 + 8 + 8 + 8 + 4 * 4 + 4 * 4) + 72;  // + NT*BN centroid norms, added at launch
}

// optional timing of the main kernel alone (bench.py's roofline leg): events are
// recorded on the launching stream around the kernel when profiling is on
static int g_ipe_force_replay = 0;
static int g_ipe_prune = 1;
static int g_filter_on = 1;           // exact / window E-step through the bf16 filter (qmeans_filter.cuh) when the shape allows
static bool g_prof_on = false;
static cudaEvent_t g_prof_ev[2] = {nullptr, nullptr};
static double g_prof_ms = 0.0;
static long long g_prof_launches = 0;

template <class Cfg>
static int launch_assign(AssignArgs a, cudaStream_t stream) {
    a.num_tiles = (int)((a.n + Cfg::BM - 1) / Cfg::BM);
    a.NT = (a.k + Cfg::BN - 1) / Cfg::BN;
    a.KC = (a.d + Cfg::BK - 1) / Cfg::BK;
    // a CTA counts its (tile, centroid block, k chunk) steps in an int: even one CTA owning every tile must fit
    if ((long long)a.num_tiles * a.NT * a.KC >= 2147483647LL) {
        set_error("sq_qmeans_assign: n=%lld d=%d k=%d is %lld tile steps, beyond 2^31 - 1; split the rows", a.n, a.d, a.k,
                  (long long)a.num_tiles * a.NT * a.KC);
        return SQ_ERR_UNSUPPORTED;
    }
    size_t smem = assign_smem_bytes<Cfg>() + (size_t)a.NT * Cfg::BN * sizeof(double);
    if (smem > 227 * 1024) { set_error("sq_qmeans_assign: k=%d needs %zu bytes of shared memory", a.k, smem); return SQ_ERR_UNSUPPORTED; }
    bool vec16 = (a.d & 1) == 0 && ((uintptr_t)a.X & 15) == 0 && ((uintptr_t)a.C & 15) == 0;
    if (g_prof_on) {
        if (!g_prof_ev[0]) { SQ_CUDA(cudaEventCreate(&g_prof_ev[0])); SQ_CUDA(cudaEventCreate(&g_prof_ev[1])); }
        SQ_CUDA(cudaEventRecord(g_prof_ev[0], stream));
    }
    // TMA path: tensor maps over X [n, d] and C [k, d] with a box of {4 doubles, rows}
    CUtensorMap tmA, tmB;
    memset(&tmA, 0, sizeof(tmA));
    memset(&tmB, 0, sizeof(tmB));
    // (SQ_ASSIGN_NO_TMA forces the cp.async path, for A/B timing; read once, when the library is loaded)
    static const bool no_tma = getenv("SQ_ASSIGN_NO_TMA") != nullptr;
    bool use_tma = vec16 && !no_tma;
    if (use_tma) {
        int r1 = make_tensor_map_2d(&tmA, a.X, (unsigned long long)a.d, (unsigned long long)a.n, 4, Cfg::BM);
        int r2 = make_tensor_map_2d(&tmB, a.C, (unsigned long long)a.d, (unsigned long long)a.k, 4, Cfg::BN);
        if (r1 || r2) use_tma = false;  // driver without the encoder: 8-byte cp.async path
    }
    smem += Cfg::STAGES * sizeof(unsigned long long) + 8 + (size_t)Cfg::BM * Cfg::BN / 8;   // barriers, two counters, IPE's keep bits
    // persistent grid: one CTA per resident slot, tiles handed out by static striding
#define SQ_LAUNCH_ASSIGN(V, M, T)                                                                                     \
    do {                                                                                                            \
        SQ_CUDA(cudaFuncSetAttribute(qmeans_assign_kernel<Cfg, V, M, T>, cudaFuncAttributeMaxDynamicSharedMemorySize, \
                                     (int)smem));                                                                   \
        int per_sm = 1;                                                                                             \
        SQ_CUDA(cudaOccupancyMaxActiveBlocksPerMultiprocessor(&per_sm, qmeans_assign_kernel<Cfg, V, M, T>,          \
                                                              Cfg::THREADS, smem));                                 \
        if (per_sm < 1) per_sm = 1;                                                                                 \
        int grid = device_sm_count() * per_sm;                                                                      \
        if (grid > a.num_tiles) grid = a.num_tiles;                                                                 \
        if (M == MODE_IPE) {                                                                                        \
            size_t fit = a.ipe_scratch_bytes / ((size_t)Cfg::BM * Cfg::BN * 12);                                    \
            if (fit < 1 || !a.ipe_scratch) {                                                                        \
                set_error("sq_qmeans_assign: ipe mode needs %zu more workspace bytes per resident CTA",             \
                          (size_t)Cfg::BM * Cfg::BN * 12);                                                          \
                return SQ_ERR_WORKSPACE;                                                                            \
            }                                                                                                       \
            if ((size_t)grid > fit) grid = (int)fit;                                                                \
        }                                                                                                           \
        qmeans_assign_kernel<Cfg, V, M, T><<<grid, Cfg::THREADS, smem, stream>>>(a, tmA, tmB);                      \
    } while (0)
    if (use_tma) {
        if (a.mode == MODE_EXACT) SQ_LAUNCH_ASSIGN(true, MODE_EXACT, true);
        else if (a.mode == MODE_WINDOW) SQ_LAUNCH_ASSIGN(true, MODE_WINDOW, true);
        else if (a.mode == MODE_IPE) SQ_LAUNCH_ASSIGN(true, MODE_IPE, true);
        else if (a.mode == MODE_COLABS) SQ_LAUNCH_ASSIGN(true, MODE_COLABS, true);
        else SQ_LAUNCH_ASSIGN(true, MODE_STORE, true);
    } else {
        if (a.mode == MODE_EXACT) SQ_LAUNCH_ASSIGN(false, MODE_EXACT, false);
        else if (a.mode == MODE_WINDOW) SQ_LAUNCH_ASSIGN(false, MODE_WINDOW, false);
        else if (a.mode == MODE_IPE) SQ_LAUNCH_ASSIGN(false, MODE_IPE, false);
        else if (a.mode == MODE_COLABS) SQ_LAUNCH_ASSIGN(false, MODE_COLABS, false);
        else SQ_LAUNCH_ASSIGN(false, MODE_STORE, false);
    }
#undef SQ_LAUNCH_ASSIGN
    if (g_prof_on) {
        // blocking on purpose: profiling mode trades overlap for a per-launch duration
        SQ_CUDA(cudaEventRecord(g_prof_ev[1], stream));
        SQ_CUDA(cudaEventSynchronize(g_prof_ev[1]));
        float ms = 0.f;
        SQ_CUDA(cudaEventElapsedTime(&ms, g_prof_ev[0], g_prof_ev[1]));
        g_prof_ms += ms;
        ++g_prof_launches;
    }
    SQ_LAUNCH_CHECK("qmeans_assign_kernel");
    return SQ_OK;
}

// tile shapes: wide (all of k <= 256 in one sweep) down to narrow for tiny k
// Two CTAs share an SM so that one CTA's epilogue (tensor pipe idle) overlaps the other's main loop.
using CfgN256 = TileCfg<32, 256, 1, 8, 2, 16, 2>;  // 256 threads, 2 x 36 KB stages, 2 CTAs/SM
using CfgN128 = TileCfg<64, 128, 2, 4, 3, 16, 2>;  // 256 threads, 3 x 24 KB stages
using CfgN64 = TileCfg<128, 64, 4, 2, 4, 16, 2>;   // 256 threads, 4 x 24 KB ring
using CfgN32 = TileCfg<128, 32, 4, 1, 4, 16, 2>;   // 128 threads
using CfgN16 = TileCfg<256, 16, 8, 1, 3, 16, 2>;   // 256 threads, k <= 16

static int dispatch_assign(const AssignArgs& a, cudaStream_t stream, int* bn_used) {
    int k = a.k;
    if (k <= 16) { *bn_used = 16; return launch_assign<CfgN16>(a, stream); }
    if (k <= 32) { *bn_used = 32; return launch_assign<CfgN32>(a, stream); }
    if (k <= 64) { *bn_used = 64; return launch_assign<CfgN64>(a, stream); }
    if (k <= 128) { *bn_used = 128; return launch_assign<CfgN128>(a, stream); }
    *bn_used = 256;
    return launch_assign<CfgN256>(a, stream);
}

}  // namespace sq

using namespace sq;

extern "C" int sq_row_sqnorms(const double* X, long long n, int d, double* out, void* stream_) {
    cudaStream_t stream = (cudaStream_t)stream_;
    SQ_REQUIRE(X && out && n >= 0 && d >= 1, "sq_row_sqnorms: bad arguments");
    if (n == 0) return SQ_OK;
    long long blocks = (n + 7) / 8;
    int cap = device_sm_count() * 16;
    if (blocks > cap) blocks = cap;
    row_sqnorm_kernel<<<(int)blocks, 256, 0, stream>>>(X, n, d, out, nullptr);
    SQ_LAUNCH_CHECK("row_sqnorm_kernel");
    return SQ_OK;
}

extern "C" int sq_sum_f64(const double* v, long long n, double* out, void* ws, size_t ws_bytes, void* stream_) {
    cudaStream_t stream = (cudaStream_t)stream_;
    SQ_REQUIRE(v && out && n >= 0, "sq_sum_f64: bad arguments");
    if (ws_bytes < SUM_BLOCKS * sizeof(double) || !ws) { set_error("sq_sum_f64: workspace needs %d bytes", (int)(SUM_BLOCKS * sizeof(double))); return SQ_ERR_WORKSPACE; }
    sum_partial_kernel<<<SUM_BLOCKS, 256, 0, stream>>>(v, n, (double*)ws);
    SQ_LAUNCH_CHECK("sum_partial_kernel");
    sum_final_kernel<<<1, 1024, 0, stream>>>((double*)ws, SUM_BLOCKS, out, nullptr, nullptr);
    SQ_LAUNCH_CHECK("sum_final_kernel");
    return SQ_OK;
}

extern "C" size_t sq_qmeans_assign_workspace_bytes(long long n, int d, int k) {
    (void)d;
    return assign_ws_layout(n, k, nullptr, nullptr) + IPE_SCRATCH_BYTES;
}

extern "C" int sq_qmeans_assign(const double* X, const double* xnorm, long long n, int d, const double* C, int k,
                                int mode, double delta, unsigned long long seed, unsigned long long offset,
                                int* labels, double* mindist, double* inertia, int* n_fixed, void* ws,
                                size_t ws_bytes, void* stream_) {
    cudaStream_t stream = (cudaStream_t)stream_;
    SQ_REQUIRE(X && xnorm && C && labels && inertia, "sq_qmeans_assign: null pointer");
    SQ_REQUIRE(n >= 1 && n < 2147483647LL && d >= 1 && k >= 1, "sq_qmeans_assign: bad shape n=%lld d=%d k=%d", n, d, k);
    SQ_REQUIRE(mode >= 0 && mode <= 2, "sq_qmeans_assign: mode must be 0 (exact), 1 (delta window) or 2 (ipe)");
    SQ_REQUIRE(mode == MODE_EXACT || delta > 0.0, "sq_qmeans_assign: noisy modes need delta > 0");
    size_t need = assign_ws_layout(n, k, nullptr, nullptr);
    if (!ws || ws_bytes < need) { set_error("sq_qmeans_assign: workspace needs %zu bytes, got %zu", need, ws_bytes); return SQ_ERR_WORKSPACE; }
    AssignWs w;
    assign_ws_layout(n, k, &w, (char*)ws);

    // centre norms; the same launch clears the near-tie counter
    {
        long long blocks = ((long long)k + 7) / 8;
        int cap = device_sm_count() * 16;
        if (blocks > cap) blocks = cap;
        row_sqnorm_kernel<<<(int)blocks, 256, 0, stream>>>(C, k, d, w.cnorm, w.flag_count);
        SQ_LAUNCH_CHECK("row_sqnorm_kernel");
    }
    int rc;

    AssignArgs a;
    a.X = X; a.xnorm = xnorm; a.C = C; a.cnorm = w.cnorm; a.rowmin_in = nullptr;
    a.labels = labels; a.mindist = mindist ? mindist : w.mindist; a.rowmin_out = nullptr;
    a.flag_count = w.flag_count; a.flag_rows = w.flag_rows;
    a.n = n; a.d = d; a.k = k; a.mode = mode; a.delta = delta;
    a.tie_tol = 16.0 * (d + 8) * 1.1102230246251565e-16;
    a.seed = seed; a.offset = offset;
    a.store = nullptr; a.colkey = nullptr;
    a.ipe_scratch = (double*)((char*)ws + need);            // whatever lies behind the fixed part
    a.ipe_scratch_bytes = ws_bytes - need;
    a.ipe_force_replay = g_ipe_force_replay;
    a.ipe_prune = g_ipe_prune;
    int bn = 0;
    if (mode == MODE_WINDOW && k > 256) {
        // the window needs the row minimum over all k before candidates can be drawn
        AssignArgs pre = a;
        pre.mode = MODE_EXACT; pre.rowmin_out = w.rowmin; pre.tie_tol = -1.0;
        rc = dispatch_assign(pre, stream, &bn);
        if (rc) return rc;
        a.rowmin_in = w.rowmin;
    }
    a.fix_mindist = 0;
    // exact and delta-window modes, all centroids in one tile: filter on the bf16 tensor cores, decide in fp64
    // (qmeans_filter.cuh); otherwise the fp64 tensor-core kernel
    const bool filter = g_filter_on && mode != MODE_IPE && k <= FilterCfg::BN && d <= FilterCfg::DMAX && d % 2 == 0 &&
                        ((uintptr_t)X & 15) == 0 && ((uintptr_t)C & 15) == 0;
    if (filter) {
        a.fix_mindist = 1;
        const size_t smem = filter_smem_bytes(d);
        long long tiles = (n + FilterCfg::BM - 1) / FilterCfg::BM;
        int grid = device_sm_count();
        if (grid > tiles) grid = (int)tiles;
        if (g_prof_on) {
            if (!g_prof_ev[0]) { SQ_CUDA(cudaEventCreate(&g_prof_ev[0])); SQ_CUDA(cudaEventCreate(&g_prof_ev[1])); }
            SQ_CUDA(cudaEventRecord(g_prof_ev[0], stream));
        }
        if (mode == MODE_EXACT) {
            SQ_CUDA(cudaFuncSetAttribute(qmeans_filter_kernel<MODE_EXACT>, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem));
            qmeans_filter_kernel<MODE_EXACT><<<grid, FilterCfg::THREADS, smem, stream>>>(a);
        } else {
            SQ_CUDA(cudaFuncSetAttribute(qmeans_filter_kernel<MODE_WINDOW>, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem));
            qmeans_filter_kernel<MODE_WINDOW><<<grid, FilterCfg::THREADS, smem, stream>>>(a);
        }
        SQ_LAUNCH_CHECK("qmeans_filter_kernel");
        if (g_prof_on) {
            cudaEventRecord(g_prof_ev[1], stream);
            cudaEventSynchronize(g_prof_ev[1]);
            float ms = 0.f;
            cudaEventElapsedTime(&ms, g_prof_ev[0], g_prof_ev[1]);
            g_prof_ms += ms;
            ++g_prof_launches;
        }
    } else {
        rc = dispatch_assign(a, stream, &bn);
        if (rc) return rc;
    }
    if (mode == MODE_EXACT) {
        int fb = device_sm_count() * 4;
        qmeans_fixup_kernel<<<fb, 256, 0, stream>>>(a);
        SQ_LAUNCH_CHECK("qmeans_fixup_kernel");
    } else if (mode == MODE_WINDOW) {
        int fb = device_sm_count() * 4;
        qmeans_window_fixup_kernel<<<fb, 256, 0, stream>>>(a);
        SQ_LAUNCH_CHECK("qmeans_window_fixup_kernel");
    }
    // inertia = sum of the row minima (fixed order); the last launch also hands out the count of re-evaluated rows
    sum_partial_kernel<<<SUM_BLOCKS, 256, 0, stream>>>(a.mindist, n, w.partial);
    SQ_LAUNCH_CHECK("sum_partial_kernel");
    sum_final_kernel<<<1, 1024, 0, stream>>>(w.partial, SUM_BLOCKS, inertia, w.flag_count, n_fixed);
    SQ_LAUNCH_CHECK("sum_final_kernel");
    return SQ_OK;
}

namespace sq {
__global__ void ipe_tail_test_kernel(double M, double x, long long B, unsigned long long seed, double* __restrict__ out) {
    long long i = (long long)blockIdx.x * blockDim.x + threadIdx.x;
    if (i >= B) return;
    const double j0 = floor(x), t = x - j0;
    out[i] = ipe_tail_draw(M, t, j0, seed, (unsigned long long)i, 0u, 0);
}
}  // namespace sq

extern "C" int sq_ipe_debug_tail_sample(double M, double x, long long B, unsigned long long seed, double* out_j,
                                        void* stream_) {
    cudaStream_t stream = (cudaStream_t)stream_;
    SQ_REQUIRE(out_j && B >= 0 && M >= 10.0 && x >= 0.0 && x < M && x != floor(x),
               "sq_ipe_debug_tail_sample: needs M >= 10 and a non-integer grid position 0 <= x < M");
    if (B == 0) return SQ_OK;
    ipe_tail_test_kernel<<<(unsigned)((B + 127) / 128), 128, 0, stream>>>(M, x, B, seed, out_j);
    SQ_LAUNCH_CHECK("ipe_tail_test_kernel");
    return SQ_OK;
}

extern "C" int sq_qmeans_debug_filter(int on) {
    g_filter_on = on ? 1 : 0;
    return SQ_OK;
}

extern "C" int sq_ipe_debug_force_replay(int on) {
    g_ipe_force_replay = (on == 1) ? 1 : 0;
    g_ipe_prune = (on == 0) ? 1 : 0;
    return SQ_OK;
}

extern "C" int sq_profile_assign(int enable, double* total_ms, long long* launches) {
    if (total_ms) *total_ms = g_prof_ms;
    if (launches) *launches = g_prof_launches;
    g_prof_on = enable != 0;
    if (enable == 2) { g_prof_ms = 0.0; g_prof_launches = 0; g_prof_on = true; }
    return SQ_OK;
}

// ---------------------------------------------------------------------------
// E-step for callers that must not allocate device memory
// ---------------------------------------------------------------------------
// Some hosts of this library cannot hand it a workspace: a plugin running
// inside a process whose device memory is entirely owned by another
// framework's pool, a latency-critical serving loop where cudaMalloc's
// implicit device-wide synchronisation is not acceptable, or glue code that
// only ever sees the caller's input and output buffers.  For them the scratch
// of sq_qmeans_assign lives in the fixed array below.  The array is part of
// the library image: every GPU that loads libsqb200.so carries these 2 MiB
// until the process exits, whether or not this entry point is ever called.
#define SQ_NOALLOC_MAX_ROWS 65536
#define SQ_NOALLOC_MAX_CLUSTERS 1024
#define SQ_NOALLOC_SCRATCH_BYTES (2u << 20)
static __device__ __align__(256) unsigned char g_noalloc_scratch[SQ_NOALLOC_SCRATCH_BYTES];

// In host code `g_noalloc_scratch` is only a handle for the symbol, not an
// address a kernel can use; the per-device address comes from the runtime.
static void* g_noalloc_ptr[64] = {nullptr};

static int noalloc_scratch_for_current_device(void** out) {
    int dev = -1;
    SQ_CUDA(cudaGetDevice(&dev));
    SQ_REQUIRE(dev >= 0 && dev < 64, "sq_qmeans_assign_noalloc: device ordinal %d not supported", dev);
    if (!g_noalloc_ptr[dev]) {
        void* p = nullptr;
        SQ_CUDA(cudaGetSymbolAddress(&p, g_noalloc_scratch));
        g_noalloc_ptr[dev] = p;
    }
    *out = g_noalloc_ptr[dev];
    return SQ_OK;
}

extern "C" int sq_qmeans_assign_noalloc(const double* X, const double* xnorm, long long n, int d, const double* C,
                                        int k, int mode, double delta, unsigned long long seed,
                                        unsigned long long offset, int* labels, double* mindist, double* inertia,
                                        int* n_fixed, void* stream_) {
    if (n > SQ_NOALLOC_MAX_ROWS || k > SQ_NOALLOC_MAX_CLUSTERS) {
        set_error("sq_qmeans_assign_noalloc: n=%lld rows / k=%d centroids exceed the built-in scratch "
                  "(at most %d rows and %d centroids per call); split the batch or use sq_qmeans_assign "
                  "with a workspace", n, k, SQ_NOALLOC_MAX_ROWS, SQ_NOALLOC_MAX_CLUSTERS);
        return SQ_ERR_UNSUPPORTED;
    }
    SQ_REQUIRE(n >= 1 && k >= 1, "sq_qmeans_assign_noalloc: bad shape n=%lld k=%d", n, k);
    size_t need = assign_ws_layout(n, k, nullptr, nullptr);
    if (need > SQ_NOALLOC_SCRATCH_BYTES) {  // cannot happen within the limits above; kept as a guard
        set_error("sq_qmeans_assign_noalloc: scratch of %zu bytes exceeds the built-in %u", need, SQ_NOALLOC_SCRATCH_BYTES);
        return SQ_ERR_UNSUPPORTED;
    }
    void* scratch = nullptr;
    int rc = noalloc_scratch_for_current_device(&scratch);
    if (rc) return rc;
    // sq_qmeans_assign zeroes the one word that has to start at zero (the near-tie counter) itself;
    // everything else in the scratch is written before it is read.
    return sq_qmeans_assign(X, xnorm, n, d, C, k, mode, delta, seed, offset, labels, mindist, inertia, n_fixed,
                            scratch, SQ_NOALLOC_SCRATCH_BYTES, stream_);
}

// ---------------------------------------------------------------------------
// K5: which row carries the largest |u_ij| of every left singular vector, and its sign
// ---------------------------------------------------------------------------
extern "C" int sq_project_absmax(const double* X, long long n, int d, const double* Vt, int r, const double* mu_v,
                                 unsigned long long* out_key, void* stream_) {
    cudaStream_t stream = (cudaStream_t)stream_;
    SQ_REQUIRE(X && Vt && mu_v && out_key, "sq_project_absmax: null pointer");
    SQ_REQUIRE(n >= 1 && n < 2147483647LL && d >= 1 && r >= 1, "sq_project_absmax: bad shape n=%lld d=%d r=%d", n, d, r);
    SQ_CUDA(cudaMemsetAsync(out_key, 0, (size_t)r * sizeof(unsigned long long), stream));
    // the columns are swept in blocks of at most 256 (one centroid tile per launch)
    for (int c0 = 0; c0 < r; c0 += 256) {
        int rc = r - c0 < 256 ? r - c0 : 256;
        AssignArgs a;
        memset(&a, 0, sizeof(a));
        a.X = X; a.C = Vt + (long long)c0 * d; a.cnorm = mu_v + c0; a.colkey = out_key + c0;
        a.n = n; a.d = d; a.k = rc; a.mode = MODE_COLABS;
        int bn = 0;
        int rcode = dispatch_assign(a, stream, &bn);
        if (rcode) return rcode;
    }
    return SQ_OK;
}

// ---------------------------------------------------------------------------
// Projection Y = X Vt^T - offset on the fp64 tensor cores (qPCA.transform)
// ---------------------------------------------------------------------------
extern "C" int sq_project(const double* X, long long n, int d, const double* Vt, int r, const double* offset,
                          double* out, void* stream_) {
    cudaStream_t stream = (cudaStream_t)stream_;
    SQ_REQUIRE(X && Vt && offset && out, "sq_project: null pointer");
    SQ_REQUIRE(n >= 1 && n < 2147483647LL && d >= 1 && r >= 1, "sq_project: bad shape n=%lld d=%d r=%d", n, d, r);
    AssignArgs a;
    memset(&a, 0, sizeof(a));
    a.X = X; a.C = Vt; a.cnorm = offset; a.store = out;
    a.n = n; a.d = d; a.k = r; a.mode = MODE_STORE;
    int bn = 0;
    return dispatch_assign(a, stream, &bn);
}
