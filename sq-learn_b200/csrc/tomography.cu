// K1 — batched vector-state tomography (L2 / L-infinity), one CTA per vector.
//
// Replaces real_tomography (sklearn/QuantumUtility/Utility.py:259-402) with
// QuantumState.measure (:51-55) and estimate_wald (:61-64) folded in.  Per
// vector V of length d and per attempt with N measurements:
//   1. V <- V/||V|| unless ||V|| is within np.isclose(.,1,rtol=1e-2) (:301-304)
//   2. stage 1: n ~ Multinomial(N, V_i^2/sum V^2), P_i = sqrt(n_i/N)   (:313,:366-369)
//   3. stage 2: 2d outcomes with amplitudes (V_i +- P_i)/2, renormalised by
//      QuantumState (:373-381, :43-47); N more draws, only "+i" counts m_i used (:385-394)
//   4. est_i = P_i if m_i > 0.4 P_i^2 N else -P_i                        (:396)
//   5. incremental mode: fresh attempts along the host-computed schedule of
//      check_measure(np.geomspace(1, N, 100)) (:317-319, :414-422); stop at the
//      first attempt whose ||V - est|| (L2 or Linf) <= delta (:355-363).
// The O(N) sampling of the reference becomes O(d): the multinomial counts are
// drawn by conditional binomial splitting down a sum tree held in shared
// memory (binomial.cuh), with Philox streams keyed by (seed, vector id,
// attempt, stage, tree node) so results do not depend on launch geometry or
// on how vectors are sharded over GPUs.
#include "binomial.cuh"
#include "common.cuh"
#include "sqb200.h"
#include "tomo_group.cuh"

namespace sq {

struct TomoArgs {
    const double* V;      // [B, d] row-major
    double* est;          // [B, d]
    double* n_used;       // [B] measurements of the returned attempt
    double* err;          // [B] ||V - est|| of the returned attempt
    double* trace;        // optional [B, sched_len, d]: the estimate of every attempt made
    int* attempts;        // optional [B]: number of attempts made
    int* stop_idx;        // [B] speculative modes: index of the first attempt that met the accuracy
    int* lane_done;       // [B] spec = 3: attempt lanes of the vector that have finished their search
    int spec;             // 0: attempts in order; 1: attempt lanes searching for stop_idx; 2: replay stop_idx;
                          // 3: search, then the lane that finishes last replays (one launch)
    double sched[128];    // attempt sizes N (integers stored as doubles), by value
    long long B;
    int d;
    int L;                // leaves padded to a power of two
    int logL;
    int sched_len;
    int norm_inf;         // 0: L2, 1: L-infinity
    int stop;             // stop at the first attempt with err <= delta
    double direct_factor; // attempts with N <= direct_factor * clamp(L, 32, 512) draw their samples one by one
    int cache;            // keep the normalised input and its stage-1 CDF in shared memory across attempts
    int early;            // give up an attempt after stage 1 when no choice of signs could meet delta
    double delta;
    unsigned long long seed;
    unsigned long long vec_offset;  // global id of vector 0 (sharding / call offset)
};

__device__ __forceinline__ double warp_scan_incl(double v) {
#pragma unroll
    for (int o = 1; o < 32; o <<= 1) {
        double y = __shfl_up_sync(0xffffffffu, v, o);
        if ((threadIdx.x & 31) >= o) v += y;
    }
    return v;
}

__device__ __forceinline__ double block_sum(double v, double* red, int tid, int nthreads) {
    v = warp_sum(v);
    __syncthreads();  // red may still be read by a previous reduction
    if ((tid & 31) == 0) red[tid >> 5] = v;
    __syncthreads();
    double t = 0.0;
    int nw = (nthreads + 31) >> 5;
    for (int w = 0; w < nw; ++w) t += red[w];  // fixed order: deterministic
    return t;
}
__device__ __forceinline__ double block_max(double v, double* red, int tid, int nthreads) {
    v = warp_max(v);
    __syncthreads();
    if ((tid & 31) == 0) red[tid >> 5] = v;
    __syncthreads();
    double t = 0.0;
    int nw = (nthreads + 31) >> 5;
    for (int w = 0; w < nw; ++w) t = fmax(t, red[w]);
    return t;
}

// bottom-up sums: mass[i] = mass[2i] + mass[2i+1]
__device__ __forceinline__ void build_tree(double* mass, int L, int logL, int tid, int T) {
    for (int lev = logL - 1; lev >= 0; --lev) {
        __syncthreads();
        int first = 1 << lev;
        for (int i = first + tid; i < 2 * first; i += T) mass[i] = mass[2 * i] + mass[2 * i + 1];
    }
    __syncthreads();
}

// Scratch of split_tree: ids of the nodes of a level whose draw binomial_head did not settle (BINV nodes from
// the front, BTRS nodes from the back), and two pairs of counters used by alternate levels.
struct SplitScratch {
    unsigned short* ids;   // [cap]
    int* ctr;              // [4]
    int cap;               // nodes of the widest level (L/2)
};
__host__ __device__ inline size_t split_scratch_doubles(int L) { return (size_t)(L / 8 > 2 ? L / 8 : 2) + 2; }
__device__ __forceinline__ SplitScratch split_scratch_at(double* mem, int L) {
    SplitScratch sc;
    sc.ctr = reinterpret_cast<int*>(mem);
    sc.ids = reinterpret_cast<unsigned short*>(mem + 2);
    sc.cap = L >> 1;
    return sc;
}

// top-down conditional binomial splitting; cnt[1] must be set by the caller.  cnt may alias mass: a node's
// two child masses are read by that node alone, right before it stores the two child counts over them.
// A level takes two passes.  Pass 1, one thread per node: binomial_head settles empty nodes and the nodes
// whose first BTRS candidate is accepted at once (most of them when n p is large); the others are listed.
// Pass 2 runs binomial_split on the listed nodes, packed densely over the threads and sorted by branch,
// so the logarithms of the full acceptance test are paid per rejected node, not per warp holding one.
// Every node draws from its own Philox stream (vec, stream, key_base + node), so neither the pass a node
// is settled in nor its place in the list changes the result.
__device__ __forceinline__ void split_tree(const double* mass, double* cnt, int logL, int tid, int T,
                                           unsigned long long seed, unsigned long long vec, unsigned stream,
                                           unsigned key_base, SplitScratch sc) {
    const int lane = tid & 31;
    if (tid == 0) { sc.ctr[0] = 0; sc.ctr[1] = 0; sc.ctr[2] = 0; sc.ctr[3] = 0; }
    for (int lev = 0; lev < logL; ++lev) {
        __syncthreads();
        const int first = 1 << lev;
        int* ctr = sc.ctr + 2 * (lev & 1);
        if (tid == 0) { sc.ctr[2 * ((lev + 1) & 1)] = 0; sc.ctr[2 * ((lev + 1) & 1) + 1] = 0; }
        for (int base = first; base < 2 * first; base += T) {
            const int i = base + tid;
            int st = BINOM_DONE;
            if (i < 2 * first) {
                const double n = cnt[i];
                double left = 0.0;
                if (n > 0.0) {
                    Philox rng(seed, vec, stream, key_base + (unsigned)i);
                    st = binomial_head(n, mass[2 * i], mass[2 * i + 1], rng, left);
                }
                if (st == BINOM_DONE) { cnt[2 * i] = left; cnt[2 * i + 1] = n - left; }
            }
            const unsigned m1 = __ballot_sync(0xffffffffu, st == BINOM_NEEDS_BINV);
            const unsigned m2 = __ballot_sync(0xffffffffu, st == BINOM_NEEDS_BTRS);
            if (m1 | m2) {
                int b1 = 0, b2 = 0;
                if (lane == 0) {
                    if (m1) b1 = atomicAdd(&ctr[0], __popc(m1));
                    if (m2) b2 = atomicAdd(&ctr[1], __popc(m2));
                }
                b1 = __shfl_sync(0xffffffffu, b1, 0);
                b2 = __shfl_sync(0xffffffffu, b2, 0);
                const unsigned lt = (1u << lane) - 1u;
                if (st == BINOM_NEEDS_BINV) sc.ids[b1 + __popc(m1 & lt)] = (unsigned short)i;
                if (st == BINOM_NEEDS_BTRS) sc.ids[sc.cap - 1 - (b2 + __popc(m2 & lt))] = (unsigned short)i;
            }
        }
        __syncthreads();
        const int nf = ctr[0], nb = ctr[1];
        for (int j = tid; j < nf + nb; j += T) {
            const int i = j < nf ? sc.ids[j] : sc.ids[sc.cap - 1 - (j - nf)];
            const double n = cnt[i];
            Philox rng(seed, vec, stream, key_base + (unsigned)i);
            const double left = binomial_split(n, mass[2 * i], mass[2 * i + 1], rng);
            cnt[2 * i] = left;
            cnt[2 * i + 1] = n - left;
        }
    }
    __syncthreads();
}

// The fused kernel's results must not depend on how many threads work on a vector (the host picks
// that from the batch size).  Draws are keyed by node / sample index already; the two helpers below
// fix the order of the floating-point additions as a function of L alone.

// sum_{i < L} f(i): pairs (i, i + L/2) are added in registers, then the L/2 partials are halved in
// scratch[0..L/2) until one is left.  f may have side effects on other arrays (it runs once per i).
template <class F>
__device__ __forceinline__ double canon_sum(F f, double* scratch, int L, int tid, int T) {
    const int h = L >> 1;
    for (int i = tid; i < h; i += T) { double lo = f(i), hi = f(i + h); scratch[i] = lo + hi; }
    int s = h >> 1;
    for (; s >= 32; s >>= 1) {
        __syncthreads();
        for (int i = tid; i < s; i += T) scratch[i] += scratch[i + s];
    }
    __syncthreads();
    // the last 2s <= 32 partials: the same pairs (i, i + s), added through shuffles by warp 0
    if (tid < 32) {
        double v = (tid < 2 * s || (s == 0 && tid == 0)) ? scratch[tid] : 0.0;
        for (; s >= 1; s >>= 1) v += __shfl_down_sync(0xffffffffu, v, s);
        if (tid == 0) scratch[0] = v;
    }
    __syncthreads();
    const double r = scratch[0];
    __syncthreads();  // scratch is reused by the caller
    return r;
}

// in-place inclusive prefix sums of x[0..L): groups of 32 consecutive entries by the warp shuffle network,
// group totals accumulated in order by warp 0 (tot: scratch of L/32 + 1 doubles)
__device__ __forceinline__ void canon_scan(double* x, int L, double* tot, int tid, int T) {
    const int lane = tid & 31, warp = tid >> 5, nw = T >> 5;
    const int G = (L + 31) >> 5;
    for (int g = warp; g < G; g += nw) {
        const int i = (g << 5) + lane;
        double s = warp_scan_incl(i < L ? x[i] : 0.0);
        if (i < L) x[i] = s;
        if (lane == 31) tot[g] = s;
    }
    __syncthreads();
    if (warp == 0) {
        double base = 0.0;
        for (int c = 0; c < G; c += 32) {
            const double t = (c + lane < G) ? tot[c + lane] : 0.0;
            const double s = warp_scan_incl(t);
            __syncwarp();
            if (c + lane < G) tot[c + lane] = base + (s - t);      // exclusive offset of the group
            base += __shfl_sync(0xffffffffu, s, 31);
        }
    }
    __syncthreads();
    for (int i = tid; i < L; i += T) x[i] += tot[i >> 5];
    __syncthreads();
}

// N independent categorical draws from the CDF cdf[0..L) with total mass tot (draws >= cdf[L-1] fall in
// the lumped remainder and are dropped); counts accumulate in cnt_i[0..L).  Sample s uses the Philox
// stream (vec, stream, s/2), so the counts do not depend on the thread count.  With enough draws to pay
// for it, a guide table (guide[j] = first bin that can hold a target in bucket j of L equal buckets) turns
// the log2(L) dependent probes of the bisection into one or two; either way the bin found is the smallest
// i with target < cdf[i].
template <bool guided>
__device__ __forceinline__ void sample_direct_impl(const double* cdf, int L, double tot, int* cnt_i, int* guide,
                                                   long long N, int tid, int T, unsigned long long seed,
                                                   unsigned long long vec, unsigned stream) {
    const double scale = (double)L / tot;
    if (guided) {
        for (int i = tid; i < L; i += T) {
            const double lo = i ? cdf[i - 1] : 0.0, hi = cdf[i];
            int j0 = (int)ceil(lo * scale), j1 = (i == L - 1) ? L : (int)ceil(hi * scale);
            j1 = min(j1, L);
            for (int j = j0; j < j1; ++j) guide[j] = i;
        }
        __syncthreads();
    }
    const long long pairs = (N + 1) >> 1;
    for (long long p = tid; p < pairs; p += T) {
        Philox rng(seed, vec, stream, (unsigned)p);
        double u[2];
        rng.uniform2(u[0], u[1]);
#pragma unroll
        for (int h = 0; h < 2; ++h) {
            if (2 * p + h >= N) break;
            const double target = u[h] * tot;
            if (target >= cdf[L - 1]) continue;
            int lo;
            if (guided) {
                lo = guide[min((int)(target * scale), L - 1)];
                while (lo > 0 && cdf[lo - 1] > target) --lo;       // (rounding of the bucket edges)
                while (cdf[lo] <= target) ++lo;
            } else {
                lo = 0;                        // smallest i with target < cdf[i]
                for (int step = L >> 1; step > 0; step >>= 1)
                    if (cdf[lo + step - 1] <= target) lo += step;
            }
            atomicAdd(&cnt_i[lo], 1);
        }
    }
}
__device__ __forceinline__ void sample_direct(const double* cdf, int L, int logL, double tot, int* cnt_i, int* guide,
                                              long long N, int tid, int T, unsigned long long seed,
                                              unsigned long long vec, unsigned stream) {
    // short vectors: the bisection is 6-8 probes, cheaper than building a table
    if (L >= 512 && N >= L) sample_direct_impl<true>(cdf, L, tot, cnt_i, guide, N, tid, T, seed, vec, stream);
    else sample_direct_impl<false>(cdf, L, tot, cnt_i, guide, N, tid, T, seed, vec, stream);
}

// shared memory of one vector, in doubles: P [L] | A [2L] (sum tree, then counts in place; CDF, int counts
// and reduction scratch on the direct path) | tot [L/32 + 1 <= 160] | split_tree's node list | with `cache`:
// X [L] (the normalised input) and C1 [L] (its stage-1 CDF), which do not change from attempt to attempt
__host__ __device__ inline size_t tomo_smem_doubles(int L, int cache) {
    return (size_t)3 * L + 160 + split_scratch_doubles(L) + (cache ? 2 * (size_t)L : 0);
}

template <int MAXT, int MINB>
__global__ void __launch_bounds__(MAXT, MINB) tomo_fused_kernel(TomoArgs a) {
    extern __shared__ double sm[];
    const int L = a.L, d = a.d, tid = threadIdx.x, T = blockDim.x, H = L >> 1;
    double* P = sm;             // [L]
    double* A = P + L;          // [2L]
    double* tot = A + 2 * L;    // [160]
    const SplitScratch sc = split_scratch_at(tot + 160, L);
    double* X = tot + 160 + split_scratch_doubles(L);   // [L], only with a.cache
    double* C1 = X + L;                                 // [L], only with a.cache
    double* mass = A;
    double* cnt = A;
    const bool cache = a.cache != 0;

    const long long b = blockIdx.x;
    const unsigned long long vec = a.vec_offset + (unsigned long long)b;
    const double* Vg = a.V + b * (long long)d;

    // The input stays in global memory (L2) and is re-read by the passes of an attempt.
    // ||V||^2 by the sum tree: an order of additions that depends on L only.
    for (int i = tid; i < L; i += T) { double x = (i < d) ? Vg[i] : 0.0; mass[L + i] = x * x; }
    build_tree(mass, L, a.logL, tid, T);
    const double nrm = sqrt(mass[1]);
    // np.isclose(nrm, 1, rtol=1e-2): |nrm - 1| <= 1e-8 + 1e-2
    const bool keep = fabs(nrm - 1.0) <= (1e-8 + 1e-2);
    if (!(nrm > 0.0)) {
        // zero (or NaN) vector: the reference divides by zero; we return zeros
        for (int i = tid; i < d; i += T) a.est[b * (long long)d + i] = 0.0;
        if (tid == 0 && a.spec != 1) { a.n_used[b] = 0.0; a.err[b] = 0.0; if (a.attempts) a.attempts[b] = 0; }
        return;
    }
    auto vin_g = [&](int i) { return (i < d) ? (keep ? Vg[i] : Vg[i] / nrm) : 0.0; };
    __syncthreads();
    double xmax2 = 0.0;
    if (cache) {
        for (int i = tid; i < L; i += T) { const double x = vin_g(i); X[i] = x; C1[i] = x * x; xmax2 = fmax(xmax2, x * x); }
        __syncthreads();
        canon_scan(C1, L, tot, tid, T);                            // stage-1 CDF, the same for every attempt
    }
    auto vin = [&](int i) { return cache ? X[i] : vin_g(i); };
    // An attempt whose stage-1 magnitudes P already miss delta for EVERY choice of signs, sum_i (|x_i| - P_i)^2 >
    // delta^2 (or max_i ||x_i| - P_i| > delta), fails whatever stage 2 draws; its stage 2 is skipped.  Attempts
    // are independent fresh draws, so this does not touch the law of the attempt that is returned - nor a bit of
    // it.  Not for the last attempt (returned even when it fails), nor when every attempt's estimate is wanted.
    const bool early_on = a.early && a.stop && a.trace == nullptr;
    const double early_thr = (a.norm_inf ? a.delta : a.delta * a.delta) * (1.0 + 1e-9);
    // The same before any draw: N measurements leave at least d - N components at zero, so the squared error
    // is at least ||x||^2 - N max_i x_i^2.
    double n_hopeless = 0.0;
    if (early_on && a.spec != 2 && cache && !a.norm_inf) {
        xmax2 = block_max(xmax2, tot, tid, T);
        n_hopeless = (C1[L - 1] - early_thr) / xmax2 * (1.0 - 1e-9);   // attempts with N < n_hopeless cannot succeed
        __syncthreads();
    }

    double n_last = 0.0, err_last = 0.0;
    int n_att = 0;
    // Speculative search (spec = 1): blockIdx.y is an attempt lane; lane g tries attempts g, g + G, ... and
    // records the smallest successful index.  Attempts are independent fresh draws keyed by their index, so
    // replaying the recorded attempt (spec = 2) returns exactly what the in-order loop would have returned.
    __shared__ int s_cut;
    int at0 = 0;
  for (int pass = 0; pass < (a.spec == 3 ? 2 : 1); ++pass) {
    const int mode = a.spec == 3 ? pass + 1 : a.spec;          // 0 in order, 1 search, 2 replay
    const bool early = early_on && mode != 2;
    int step = 1, at_end = a.sched_len;
    at0 = 0;
    if (mode == 1) { at0 = blockIdx.y; step = gridDim.y; }
    if (mode == 2) {
        __syncthreads();
        if (tid == 0) s_cut = *((volatile int*)&a.stop_idx[b]);
        __syncthreads();
        at0 = min(s_cut, a.sched_len - 1); at_end = at0 + 1;
        __syncthreads();
        if (a.spec == 3 && tid == 0) a.stop_idx[b] = 0x7fffffff;      // as the next call expects it
    }
    for (int at = at0; at < at_end; at += step) {
        if (mode == 1) {
            __syncthreads();
            if (tid == 0) s_cut = *((volatile int*)&a.stop_idx[b]);
            __syncthreads();
            if (at > s_cut) break;          // an earlier attempt already succeeded
        }
        const double N = a.sched[at];
        const bool may_skip = early && at + 1 < a.sched_len;
        if (may_skip && N < n_hopeless) { n_att = at + 1; continue; }
        // Few measurements: draw them one by one (N/T categorical draws per thread, log2(d) probes each)
        // instead of walking the tree, whose log2(d) dependent binomial draws cost more than that.
        // The rule depends on (N, L) only.
        const bool direct = N <= a.direct_factor * (L < 32 ? 32 : (L > 512 ? 512 : L));
        int* cnt_i = reinterpret_cast<int*>(A + L);      // [L] ints = A[L .. L + L/2)
        int* guide = reinterpret_cast<int*>(A + L + H);  // [L] ints = A[L + L/2 .. 2L), shared with the reduction scratch
        double e;
        if (direct) {
            const double* cdf1 = cache ? C1 : A;
            if (cache) {
                for (int i = tid; i < L; i += T) cnt_i[i] = 0;
                __syncthreads();
            } else {
                for (int i = tid; i < L; i += T) { double x = vin(i); A[i] = x * x; cnt_i[i] = 0; }
                __syncthreads();
                canon_scan(A, L, tot, tid, T);                     // stage-1 CDF in A[0..L)
            }
            sample_direct(cdf1, L, a.logL, cdf1[L - 1], cnt_i, guide, (long long)N, tid, T, a.seed, vec, (unsigned)(2 * at));
            __syncthreads();
            if (may_skip) {
                double lb = 0.0;
                for (int i = tid; i < L; i += T) {
                    const double g = fabs(vin(i)) - sqrt((double)cnt_i[i] / N);
                    lb = a.norm_inf ? fmax(lb, fabs(g)) : lb + g * g;
                }
                lb = a.norm_inf ? block_max(lb, tot, tid, T) : block_sum(lb, tot, tid, T);
                __syncthreads();
                if (lb > early_thr) { n_att = at + 1; continue; }
            }
            const double sminus = canon_sum([&](int i) {
                double p = sqrt((double)cnt_i[i] / N), x = vin(i);
                P[i] = p;
                double ap = 0.5 * (x + p), am = 0.5 * (x - p);
                A[i] = ap * ap;                                    // (the stage-1 CDF is not needed any more)
                return am * am;
            }, A + L + H, L, tid, T);
            for (int i = tid; i < L; i += T) cnt_i[i] = 0;
            __syncthreads();
            canon_scan(A, L, tot, tid, T);                         // stage-2 '+' CDF; the '-' outcomes are one lump
            sample_direct(A, L, a.logL, A[L - 1] + sminus, cnt_i, guide, (long long)N, tid, T, a.seed, vec,
                          (unsigned)(2 * at + 1));
            __syncthreads();
        } else {
            // ---- stage 1 ----
            for (int i = tid; i < L; i += T) { double x = vin(i); mass[L + i] = x * x; }
            build_tree(mass, L, a.logL, tid, T);
            if (tid == 0) cnt[1] = N;
            split_tree(mass, cnt, a.logL, tid, T, a.seed, vec, (unsigned)(2 * at), 0u, sc);
            if (may_skip) {
                double lb = 0.0;
                for (int i = tid; i < L; i += T) {
                    const double g = fabs(vin(i)) - sqrt(cnt[L + i] / N);
                    lb = a.norm_inf ? fmax(lb, fabs(g)) : lb + g * g;
                }
                lb = a.norm_inf ? block_max(lb, tot, tid, T) : block_sum(lb, tot, tid, T);
                __syncthreads();
                if (lb > early_thr) { n_att = at + 1; continue; }
            }
            // ---- stage 2 ----
            const double sminus = canon_sum([&](int i) {
                double p = sqrt(cnt[L + i] / N), x = vin(i);
                P[i] = p;
                double ap = 0.5 * (x + p), am = 0.5 * (x - p);
                mass[L + i] = ap * ap;                             // (leaf slot i: read and written by this call only)
                return am * am;
            }, A, L, tid, T);
            build_tree(mass, L, a.logL, tid, T);
            if (tid == 0) {
                Philox rng(a.seed, vec, (unsigned)(2 * at + 1), 0u);
                cnt[1] = binomial_split(N, mass[1], sminus, rng);
            }
            split_tree(mass, cnt, a.logL, tid, T, a.seed, vec, (unsigned)(2 * at + 1), 0u, sc);
        }
        // ---- sign rule + error ----
        auto sign_and_diff = [&](int i) {
            double p = P[i];
            double c2 = direct ? (double)cnt_i[i] : cnt[L + i];
            double est = (c2 > tomo_sign_threshold(p, N)) ? p : -p;
            P[i] = est;
            return (i < d) ? vin(i) - est : 0.0;
        };
        if (a.norm_inf) {
            double m = 0.0;
            for (int i = tid; i < L; i += T) m = fmax(m, fabs(sign_and_diff(i)));
            e = block_max(m, tot, tid, T);
        } else {
            e = sqrt(canon_sum([&](int i) { double df = sign_and_diff(i); return df * df; }, A, L, tid, T));
        }
        n_last = N;
        err_last = e;
        n_att = at + 1;
        if (a.trace) {
            __syncthreads();
            double* tr = a.trace + (b * a.sched_len + at) * (long long)d;
            for (int i = tid; i < d; i += T) tr[i] = P[i];
        }
        if (mode == 1) {
            if (e <= a.delta) { if (tid == 0) atomicMin(&a.stop_idx[b], at); break; }
            continue;
        }
        if (a.stop && e <= a.delta) break;  // e is identical in every thread
    }
    if (mode == 1) {
        if (a.spec != 3) return;
        // the lane that finishes last knows the first successful attempt of all lanes and replays it
        __shared__ int s_last;
        __syncthreads();
        if (tid == 0) {
            __threadfence();
            s_last = (atomicAdd(&a.lane_done[b], 1) == (int)gridDim.y - 1);
        }
        __syncthreads();
        if (!s_last) return;
        if (tid == 0) { a.lane_done[b] = 0; __threadfence(); }
    }
  }
    if (a.spec >= 2) n_att = at0 + 1;
    __syncthreads();
    for (int i = tid; i < d; i += T) a.est[b * (long long)d + i] = P[i];
    if (tid == 0) { a.n_used[b] = n_last; a.err[b] = err_last; if (a.attempts) a.attempts[b] = n_att; }
}

// Threads per vector: with few vectors in flight the latency of a vector counts (one thread per leaf, at most 512);
// with many, throughput does (L/4 threads, at most 128: few threads per vector keep the lanes busy through the upper
// levels of the tree and let 9+ vectors share an SM).  The results do not depend on the choice.
// Tunables (sq_tomography_tune; process-wide, not thread-safe: a tool for measurements, not part of the API
// a caller needs).  They change how the work is arranged, and direct_factor which of the two exact samplers
// an attempt uses - never the law of the result.
static double g_direct_factor = 64.0;
static int g_many_div = 4, g_many_cap = 128;
static int g_cache_max_L = 256;       // the fused kernel keeps input + stage-1 CDF in shared memory up to this L
static int g_early = 1;               // skip stage 2 of attempts that cannot succeed
static int g_blockgroup_on = 1;       // leaf blocks of long vectors go through the block-group kernel (tomo_group.cu)
static int g_group_on = 1;            // single attempts of many vectors go to the group kernel (tomo_group.cu)
static long long g_group_min_B = 296; // ... from this many vectors on
static int g_group_threads = 256;     // threads per CTA of the group kernel (1024 / threads CTAs per SM)

static int tomo_launch(TomoArgs a, long long B, int lanes, cudaStream_t stream) {
    const int L = a.L;
    a.direct_factor = g_direct_factor;
    const bool many = B * (long long)lanes >= 4096;
    int threads = many ? (L / g_many_div > g_many_cap ? g_many_cap : L / g_many_div) : L;   // (128 = the widest launch of the 64-register build)
    threads = (threads + 31) & ~31;
    threads = threads < 32 ? 32 : (threads > 512 ? 512 : threads);
    a.cache = (L <= g_cache_max_L && a.sched_len > 1) ? 1 : 0;
    a.early = g_early;
    const size_t smem = tomo_smem_doubles(L, a.cache) * sizeof(double);
    // up to 128 threads per vector: capped at 64 registers (a few spilled words), 8+ CTAs per SM - the kernel is
    // bound by dependent FP64 latency, and 64 registers measured best against 48, 80, 96 and 126
    if (threads <= 128) {
        SQ_CUDA(cudaFuncSetAttribute(tomo_fused_kernel<128, 8>, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem));
        tomo_fused_kernel<128, 8><<<dim3((unsigned)B, lanes), threads, smem, stream>>>(a);
    } else {
        SQ_CUDA(cudaFuncSetAttribute(tomo_fused_kernel<512, 1>, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem));
        tomo_fused_kernel<512, 1><<<dim3((unsigned)B, lanes), threads, smem, stream>>>(a);
    }
    SQ_LAUNCH_CHECK("tomo_fused_kernel");
    return SQ_OK;
}

// ---------------------------------------------------------------------------
// Long vectors (d > 4096): the same algorithm split over the grid.
// A vector is cut into blocks of TB leaves.  Per attempt: block masses -> one CTA splits N over the
// blocks (top tree) -> every block splits its share over its leaves in shared memory; twice (stage 1,
// stage 2), then the error is reduced and the stop rule applied.  Scratch lives in the caller's
// workspace; the stage-1 estimate P is parked in `est` between the two stages.
// ---------------------------------------------------------------------------
static const int TB = 1024;       // leaves per block
static const int TB_LOG = 10;

struct TomoBigArgs {
    const double* V;
    double* est;
    double* n_used;
    double* err;
    const double* sumsq;   // [B] ||V||^2
    double* bm_0;          // [B, nblk] stage-1 mass of every block (the same for every attempt: computed once)
    double* bm_a;          // [B, nblk] stage-2 '+' mass of every block
    double* bm_b;          // [B, nblk] stage-2 '-' mass
    double* bcnt;          // [B, nblk] measurements that fall in the block
    double* errp;          // [B, nblk] per-block error partial
    int* done;             // [B]
    double* nhop;          // [B] attempts with fewer measurements than this cannot meet delta (0: no such bound)
    double* lbp;           // [B, nblk] per block: sum (or max) of (|x_i| - P_i)^2 after stage 1
    int* skip2;            // [B] stage 2 of the current attempt is skipped
    int may_skip;          // the current attempt may be given up early (stop rule on, not the last attempt)
    int early;             // early give-up enabled for this call (phase 0 then computes nhop)
    double* gws;           // trees of the block-group kernel (tomo_group.cu)
    long long B;
    int d, nblk, NB, logNB;
    int blk_lo;            // first block of this launch (leaf blocks can be sharded over GPUs)
    double N;
    int at, norm_inf, stop;
    double delta;
    unsigned long long seed, vec_offset;
};

__device__ __forceinline__ double big_scale(double sumsq, bool& zero) {
    double nrm = sqrt(sumsq);
    zero = !(nrm > 0.0);
    bool keep = fabs(nrm - 1.0) <= (1e-8 + 1e-2);
    return keep ? 1.0 : nrm;   // divisor
}

// one CTA per vector: done flags, and the number of measurements below which an attempt cannot succeed: N
// measurements leave at least d - N components at zero, so the squared L2 error is at least
// ||x||^2 - N max_i x_i^2 (x the normalised input); an attempt with that above delta^2 draws nothing.
__global__ void tomo_big_init_kernel(TomoBigArgs a, int early) {
    __shared__ double red[32];
    const long long b = blockIdx.x;
    const int tid = threadIdx.x;
    double m = 0.0;
    if (early) {
        const double* v = a.V + b * (long long)a.d;
        for (int i = tid; i < a.d; i += blockDim.x) m = fmax(m, fabs(v[i]));
        m = block_max(m, red, tid, blockDim.x);
    }
    if (tid == 0) {
        const double ss = a.sumsq[b];
        a.done[b] = (ss > 0.0) ? 0 : 1;
        a.n_used[b] = 0.0;
        a.err[b] = 0.0;
        a.skip2[b] = 0;
        double nh = 0.0;
        if (early && ss > 0.0 && m > 0.0) {
            bool zero;
            const double div = big_scale(ss, zero);
            const double S = ss / (div * div), x2 = (m / div) * (m / div);
            nh = (S - a.delta * a.delta * (1.0 + 1e-9)) / x2 * (1.0 - 1e-9);
        }
        a.nhop[b] = nh;
    }
}
#define TOMO_BIG_IDLE(a, b) ((a).done[b] || ((a).may_skip && (a).N < (a).nhop[b]))

// block masses of stage 1 (and zero output for zero vectors)
__global__ void tomo_big_mass_kernel(TomoBigArgs a) {
    __shared__ double red[32];
    const long long b = blockIdx.y;
    const int blk = blockIdx.x, tid = threadIdx.x;
    bool zero;
    const double div = big_scale(a.sumsq[b], zero);
    if (zero) {
        for (int i = tid; i < TB; i += blockDim.x) { int g = blk * TB + i; if (g < a.d) a.est[b * (long long)a.d + g] = 0.0; }
        return;
    }
    if (TOMO_BIG_IDLE(a, b)) return;
    double s = 0.0;
    for (int i = tid; i < TB; i += blockDim.x) {
        int g = blk * TB + i;
        double x = (g < a.d) ? a.V[b * (long long)a.d + g] / div : 0.0;
        s += x * x;
    }
    s = block_sum(s, red, tid, blockDim.x);
    if (tid == 0) a.bm_0[b * a.nblk + blk] = s;
}

// one CTA per vector: split the attempt's measurements over the blocks
__global__ void tomo_big_top_kernel(TomoBigArgs a, int stage) {
    extern __shared__ double sm[];
    const long long b = blockIdx.x;
    if (TOMO_BIG_IDLE(a, b)) return;
    const int tid = threadIdx.x, T = blockDim.x, NB = a.NB;
    double* mass = sm;            // [2 NB]
    double* cnt = mass + 2 * NB;  // [2 NB]
    double* red = cnt + 2 * NB;   // [32]
    const SplitScratch sc = split_scratch_at(red + 32, NB);
    const unsigned long long vec = a.vec_offset + (unsigned long long)b;
    if (stage == 1 && tid == 0) a.skip2[b] = 0;
    if (stage == 2 && a.may_skip) {
        // stage 1 left P; if no choice of signs can bring ||x - est|| under delta, stage 2 is not drawn
        double lb = 0.0;
        for (int i = tid; i < a.nblk; i += T) {
            const double p = a.lbp[b * a.nblk + i];
            lb = a.norm_inf ? fmax(lb, p) : lb + p;
        }
        lb = a.norm_inf ? block_max(lb, red, tid, T) : block_sum(lb, red, tid, T);
        const bool hopeless = lb > (a.norm_inf ? a.delta : a.delta * a.delta) * (1.0 + 1e-9);
        if (tid == 0) a.skip2[b] = hopeless ? 1 : 0;
        if (hopeless) return;
        __syncthreads();
    }
    double sminus = 0.0;
    for (int i = tid; i < NB; i += T) {
        mass[NB + i] = (i < a.nblk) ? (stage == 1 ? a.bm_0 : a.bm_a)[b * a.nblk + i] : 0.0;
        if (stage == 2 && i < a.nblk) sminus += a.bm_b[b * a.nblk + i];
    }
    if (stage == 2) sminus = block_sum(sminus, red, tid, T);
    build_tree(mass, NB, a.logNB, tid, T);
    if (tid == 0) {
        if (stage == 1) cnt[1] = a.N;
        else {
            Philox rng(a.seed, vec, (unsigned)(4 * a.at + 2), 0u);
            cnt[1] = binomial_split(a.N, mass[1], sminus, rng);
        }
    }
    split_tree(mass, cnt, a.logNB, tid, T, a.seed, vec, (unsigned)(4 * a.at + (stage == 1 ? 0 : 2)), 0u, sc);
    for (int i = tid; i < a.nblk; i += T) a.bcnt[b * a.nblk + i] = cnt[NB + i];
}

// one CTA per block: split the block's share over its leaves
__global__ void tomo_big_block_kernel(TomoBigArgs a, int stage) {
    extern __shared__ double sm[];
    const long long b = blockIdx.y;
    if (TOMO_BIG_IDLE(a, b) || (stage == 2 && a.skip2[b])) return;
    const int blk = blockIdx.x + a.blk_lo, tid = threadIdx.x, T = blockDim.x;
    double* v = sm;               // [TB]
    double* P = v + TB;           // [TB]
    double* mass = P + TB;        // [2 TB]
    double* cnt = mass + 2 * TB;  // [2 TB]
    double* red = cnt + 2 * TB;   // [32]
    const SplitScratch sc = split_scratch_at(red + 32, TB);
    bool zero;
    const double div = big_scale(a.sumsq[b], zero);
    const unsigned long long vec = a.vec_offset + (unsigned long long)b;
    const long long base = b * (long long)a.d + (long long)blk * TB;
    for (int i = tid; i < TB; i += T) {
        int g = blk * TB + i;
        double x = (g < a.d) ? a.V[base + i] / div : 0.0;
        v[i] = x;
        if (stage == 1) mass[TB + i] = x * x;
        else {
            double p = (g < a.d) ? a.est[base + i] : 0.0;
            P[i] = p;
            double ap = 0.5 * (x + p);
            mass[TB + i] = ap * ap;
        }
    }
    build_tree(mass, TB, TB_LOG, tid, T);
    if (tid == 0) cnt[1] = a.bcnt[b * a.nblk + blk];
    // node ids are offset by the block so that every (vector, block, node) owns its own Philox stream
    split_tree(mass, cnt, TB_LOG, tid, T, a.seed, vec, (unsigned)(4 * a.at + (stage == 1 ? 1 : 3)),
               (unsigned)(blk * (2 * TB)), sc);
    if (stage == 1) {
        double sp = 0.0, smn = 0.0, lb = 0.0;
        for (int i = tid; i < TB; i += T) {
            int g = blk * TB + i;
            double p = sqrt(cnt[TB + i] / a.N);
            if (g < a.d) a.est[base + i] = p;
            double ap = 0.5 * (v[i] + p), am = 0.5 * (v[i] - p);
            sp += ap * ap;
            smn += am * am;
            const double gap = fabs(v[i]) - p;
            lb = a.norm_inf ? fmax(lb, fabs(gap)) : lb + gap * gap;
        }
        sp = block_sum(sp, red, tid, T);
        smn = block_sum(smn, red, tid, T);
        if (a.may_skip) lb = a.norm_inf ? block_max(lb, red, tid, T) : block_sum(lb, red, tid, T);
        if (tid == 0) {
            a.bm_a[b * a.nblk + blk] = sp; a.bm_b[b * a.nblk + blk] = smn;
            if (a.may_skip) a.lbp[b * a.nblk + blk] = lb;
        }
    } else {
        double e = 0.0;
        for (int i = tid; i < TB; i += T) {
            int g = blk * TB + i;
            double p = P[i];
            double est = (cnt[TB + i] > tomo_sign_threshold(p, a.N)) ? p : -p;
            if (g < a.d) {
                a.est[base + i] = est;
                double df = v[i] - est;
                e = a.norm_inf ? fmax(e, fabs(df)) : e + df * df;
            }
        }
        e = a.norm_inf ? block_max(e, red, tid, T) : block_sum(e, red, tid, T);
        if (tid == 0) a.errp[b * a.nblk + blk] = e;
    }
}

__global__ void tomo_big_finish_kernel(TomoBigArgs a) {
    __shared__ double red[32];
    const long long b = blockIdx.x;
    if (TOMO_BIG_IDLE(a, b) || a.skip2[b]) return;
    const int tid = threadIdx.x;
    double e = 0.0;
    for (int i = tid; i < a.nblk; i += blockDim.x) {
        double p = a.errp[b * a.nblk + i];
        e = a.norm_inf ? fmax(e, p) : e + p;
    }
    e = a.norm_inf ? block_max(e, red, tid, blockDim.x) : sqrt(block_sum(e, red, tid, blockDim.x));
    if (tid == 0) {
        a.n_used[b] = a.N;
        a.err[b] = e;
        if (a.stop && e <= a.delta) a.done[b] = 1;
    }
}

}  // namespace sq

// ---------------------------------------------------------------------------
// C ABI
// ---------------------------------------------------------------------------
using namespace sq;

extern "C" int sq_tomography_max_fused_dim(void) { return 4096; }

extern "C" int sq_tomography_tune(int what, double value) {
    switch (what) {
        case 0: SQ_REQUIRE(value >= 0.0, "sq_tomography_tune: direct_factor >= 0"); g_direct_factor = value; break;
        case 1: SQ_REQUIRE(value >= 1.0 && value <= 64.0, "sq_tomography_tune: divisor in [1, 64]"); g_many_div = (int)value; break;
        case 2: SQ_REQUIRE(value >= 32.0 && value <= 128.0, "sq_tomography_tune: cap in [32, 128]"); g_many_cap = (int)value; break;
        case 3: g_group_on = value != 0.0; break;
        case 6: g_cache_max_L = (int)value; break;
        case 8: g_blockgroup_on = value != 0.0; break;
        case 7: g_early = value != 0.0; break;
        case 5: SQ_REQUIRE(value == 256.0, "sq_tomography_tune: the group kernel is built for 256 threads per CTA");
                g_group_threads = (int)value; break;
        case 4: SQ_REQUIRE(value >= 1.0, "sq_tomography_tune: minimum batch >= 1"); g_group_min_B = (long long)value; break;
        default: SQ_REQUIRE(false, "sq_tomography_tune: unknown tunable %d", what);
    }
    return SQ_OK;
}

extern "C" int sq_tomography_batched(const double* V, long long B, int d, double delta, int norm_inf, int stop,
                                     const double* sched_host, int sched_len, unsigned long long seed,
                                     unsigned long long vec_offset, double* est, double* n_used, double* err,
                                     double* trace, int* attempts, void* stream_) {
    cudaStream_t stream = (cudaStream_t)stream_;
    SQ_REQUIRE(V && est && n_used && err && sched_host, "sq_tomography_batched: null pointer");
    SQ_REQUIRE(B >= 0 && d >= 1, "sq_tomography_batched: bad shape B=%lld d=%d", B, d);
    SQ_REQUIRE(sched_len >= 1 && sched_len <= 128, "sq_tomography_batched: sched_len=%d outside [1,128]", sched_len);
    SQ_REQUIRE(delta > 0.0, "sq_tomography_batched: delta must be > 0");
    if (d > sq_tomography_max_fused_dim()) {
        set_error("sq_tomography_batched: d=%d exceeds the fused-path limit %d", d, sq_tomography_max_fused_dim());
        return SQ_ERR_UNSUPPORTED;
    }
    if (B == 0) return SQ_OK;
    TomoArgs a;
    a.V = V; a.est = est; a.n_used = n_used; a.err = err; a.trace = trace; a.attempts = attempts;
    a.stop_idx = nullptr; a.lane_done = nullptr; a.spec = 0;
    for (int i = 0; i < sched_len; ++i) {
        SQ_REQUIRE(sched_host[i] >= 1.0 && sched_host[i] < 9.0e15, "sq_tomography_batched: schedule entry out of range");
        a.sched[i] = sched_host[i];
    }
    a.B = B; a.d = d;
    int L = 2, logL = 1;
    while (L < d) { L <<= 1; ++logL; }
    a.L = L; a.logL = logL; a.sched_len = sched_len; a.norm_inf = norm_inf; a.stop = stop; a.delta = delta;
    a.seed = seed; a.vec_offset = vec_offset;
    // gridDim.x limit is 2^31-1; batches beyond that are split
    long long done = 0;
    while (done < B) {
        long long nb = B - done > 1000000000LL ? 1000000000LL : B - done;
        TomoArgs c = a;
        c.V = V + done * d; c.est = est + done * d; c.n_used = n_used + done; c.err = err + done;
        if (trace) c.trace = trace + done * sched_len * d;
        if (attempts) c.attempts = attempts + done;
        c.vec_offset = vec_offset + (unsigned long long)done;
        c.B = nb;
        int rc = tomo_launch(c, nb, 1, stream);
        if (rc) return rc;
        done += nb;
    }
    return SQ_OK;
}

// ---- long vectors ------------------------------------------------------------
extern "C" int sq_row_sqnorms(const double* X, long long n, int d, double* out, void* stream);

static size_t tomo_blockgroup_slots(long long members) {
    const long long most = (long long)device_sm_count() * 4 * 8;     // 4 CTAs per SM, up to 8 members each
    return (size_t)(members + 8 < most ? members + 8 : most);
}

static void tomo_big_layout(long long B, int d, int* nblk, int* NB, int* logNB, size_t off[11], size_t* total) {
    int nb = (d + TB - 1) / TB;
    int p2 = 2, lg = 1;
    while (p2 < nb) { p2 <<= 1; ++lg; }
    *nblk = nb; *NB = p2; *logNB = lg;
    size_t o = 0;
    size_t per = align_up((size_t)B * nb * sizeof(double), 256);
    off[0] = o; o += align_up((size_t)B * sizeof(double), 256);   // sumsq
    off[1] = o; o += per;                                         // bm_a
    off[2] = o; o += per;                                         // bm_b
    off[3] = o; o += per;                                         // bcnt
    off[4] = o; o += per;                                         // errp
    off[5] = o; o += align_up((size_t)B * sizeof(int), 256);      // done
    off[6] = o; o += align_up((size_t)B * sizeof(double), 256);   // nhop
    off[7] = o; o += per;                                         // lbp
    off[8] = o; o += align_up((size_t)B * sizeof(int), 256);      // skip2
    off[9] = o; o += tomo_blockgroup_slots(B * (long long)nb) * 2048 * sizeof(double);   // trees of the block-group kernel
    off[10] = o; o += per;                                        // bm_0
    *total = o;
}

// group kernel (d <= the fused limit): geometry for a batch of B vectors of L leaves
static void tomo_group_geometry(long long B, int L, int* G, int* ctas) {
    const int sms = device_sm_count();
    *G = tomo_group_size(B, L, sms, g_group_threads);
    const long long groups = (B + *G - 1) / *G, most = (long long)sms * (1024 / g_group_threads);
    *ctas = (int)(groups < most ? groups : most);
}
static void tomo_leaves(int d, int* L, int* logL) {
    int l = 2, lg = 1;
    while (l < d) { l <<= 1; ++lg; }
    *L = l; *logL = lg;
}

extern "C" size_t sq_tomography_workspace_bytes(long long B, int d) {
    if (B <= 0 || d < 1) return 0;
    if (d <= sq_tomography_max_fused_dim()) {
        if (!g_group_on || B < g_group_min_B) return 0;
        int L, logL, G, ctas;
        tomo_leaves(d, &L, &logL);
        tomo_group_geometry(B, L, &G, &ctas);
        return tomo_group_ws_bytes(L, G, ctas);
    }
    int nblk, NB, logNB;
    size_t off[11], total;
    tomo_big_layout(B, d, &nblk, &NB, &logNB, off, &total);
    return total;
}

// phases of one attempt on the long-vector path; [blk_lo, blk_hi) is the range of leaf blocks this call splits
static int tomo_big_prepare(TomoBigArgs& a, const double* V, long long B, int d, double delta, int norm_inf, int stop,
                            unsigned long long seed, unsigned long long vec_offset, double* est, double* n_used,
                            double* err, void* ws, size_t ws_bytes, const char* who) {
    int nblk, NB, logNB;
    size_t off[11], total;
    tomo_big_layout(B, d, &nblk, &NB, &logNB, off, &total);
    if (NB > 4096) { set_error("%s: d=%d exceeds the supported %d", who, d, 4096 * TB); return SQ_ERR_UNSUPPORTED; }
    if (!ws || ws_bytes < total) { set_error("%s: workspace needs %zu bytes, got %zu", who, total, ws_bytes); return SQ_ERR_WORKSPACE; }
    char* w = (char*)ws;
    a.V = V; a.est = est; a.n_used = n_used; a.err = err;
    a.sumsq = (double*)(w + off[0]); a.bm_a = (double*)(w + off[1]); a.bm_b = (double*)(w + off[2]);
    a.bcnt = (double*)(w + off[3]); a.errp = (double*)(w + off[4]); a.done = (int*)(w + off[5]);
    a.nhop = (double*)(w + off[6]); a.lbp = (double*)(w + off[7]); a.skip2 = (int*)(w + off[8]); a.may_skip = 0; a.early = 0;
    a.gws = (double*)(w + off[9]); a.bm_0 = (double*)(w + off[10]);
    a.B = B; a.d = d; a.nblk = nblk; a.NB = NB; a.logNB = logNB; a.norm_inf = norm_inf; a.stop = stop; a.delta = delta;
    a.seed = seed; a.vec_offset = vec_offset; a.blk_lo = 0; a.at = 0; a.N = 1.0;
    return SQ_OK;
}

static int tomo_big_blocks(const TomoBigArgs& a, int stage, int blk_lo, int blk_hi, dim3 gown, size_t sm_blk, cudaStream_t stream) {
    if (!g_blockgroup_on) {
        tomo_big_block_kernel<<<gown, 256, sm_blk, stream>>>(a, stage);
        return SQ_OK;
    }
    TomoBlockGroupArgs g;
    g.V = a.V; g.est = a.est; g.sumsq = a.sumsq; g.bcnt = a.bcnt; g.bm_a = a.bm_a; g.bm_b = a.bm_b; g.lbp = a.lbp;
    g.errp = a.errp; g.done = a.done; g.nhop = a.nhop; g.skip2 = a.skip2; g.gws = a.gws; g.B = a.B; g.d = a.d;
    g.nblk = a.nblk; g.blk_lo = blk_lo; g.nown = blk_hi - blk_lo; g.stage = stage; g.at = a.at; g.norm_inf = a.norm_inf;
    g.may_skip = a.may_skip; g.N = a.N; g.seed = a.seed; g.vec_offset = a.vec_offset;
    const long long members = a.B * (long long)g.nown, most = (long long)device_sm_count() * 4;
    long long G = (members + most - 1) / most;
    G = G < 1 ? 1 : (G > 8 ? 8 : G);
    long long ctas = (members + G - 1) / G;
    if (ctas > most) ctas = most;
    g.G = (int)G;
    return tomo_blockgroup_launch(g, (int)ctas, stream);
}

static int tomo_big_phase(TomoBigArgs a, int phase, int blk_lo, int blk_hi, cudaStream_t stream) {
    const size_t sm_top = (size_t)(4 * a.NB + 32 + split_scratch_doubles(a.NB)) * sizeof(double);
    const size_t sm_blk = (size_t)(6 * TB + 32 + split_scratch_doubles(TB)) * sizeof(double);
    const int ttop = a.NB < 64 ? 64 : (a.NB > 512 ? 512 : a.NB);   // register budget: ~96 per thread
    const dim3 gall(a.nblk, (unsigned)a.B), gown(blk_hi - blk_lo, (unsigned)a.B);
    a.blk_lo = blk_lo;
    if (phase == 0) {
        int rc = sq_row_sqnorms(a.V, a.B, a.d, (double*)a.sumsq, stream);
        if (rc) return rc;
        tomo_big_init_kernel<<<(unsigned)a.B, 256, 0, stream>>>(a, a.early && !a.norm_inf);
        tomo_big_mass_kernel<<<gall, 256, 0, stream>>>(a);       // stage-1 block masses: once, every block
        SQ_CUDA(cudaFuncSetAttribute(tomo_big_top_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)sm_top));
        SQ_CUDA(cudaFuncSetAttribute(tomo_big_block_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)sm_blk));
    } else if (phase == 1) {
        // stage 1: top tree over the block masses of phase 0, leaves of the own blocks
        tomo_big_top_kernel<<<(unsigned)a.B, ttop, sm_top, stream>>>(a, 1);
        if (blk_hi > blk_lo) { int rc = tomo_big_blocks(a, 1, blk_lo, blk_hi, gown, sm_blk, stream); if (rc) return rc; }
    } else if (phase == 2) {
        // stage 2 (needs the '+' / '-' masses of every block)
        tomo_big_top_kernel<<<(unsigned)a.B, ttop, sm_top, stream>>>(a, 2);
        if (blk_hi > blk_lo) { int rc = tomo_big_blocks(a, 2, blk_lo, blk_hi, gown, sm_blk, stream); if (rc) return rc; }
    } else {
        // error over all blocks, stop rule
        tomo_big_finish_kernel<<<(unsigned)a.B, 256, 0, stream>>>(a);
    }
    SQ_LAUNCH_CHECK("tomo_big kernels");
    return SQ_OK;
}

extern "C" int sq_tomography_batched_ws(const double* V, long long B, int d, double delta, int norm_inf, int stop,
                                        const double* sched_host, int sched_len, unsigned long long seed,
                                        unsigned long long vec_offset, double* est, double* n_used, double* err,
                                        void* ws, size_t ws_bytes, void* stream_) {
    cudaStream_t stream = (cudaStream_t)stream_;
    if (d <= sq_tomography_max_fused_dim()) {
        // One attempt (single-shot, or the last of a schedule without the stop rule) of many vectors, with more
        // measurements than the one-by-one sampler takes: the group kernel.  Its results are bit-identical to the
        // fused kernel's, so which of the two runs is a matter of speed only.
        const bool one = sched_len == 1 || !stop;
        if (one && g_group_on && B >= g_group_min_B && V && est && n_used && err && sched_host && ws &&
            sched_len >= 1 && sched_len <= 128 && d >= 1 && delta > 0.0) {
            int L, logL, G, ctas;
            tomo_leaves(d, &L, &logL);
            const double N = sched_host[sched_len - 1];
            SQ_REQUIRE(N >= 1.0 && N < 9.0e15, "sq_tomography_batched_ws: schedule entry out of range");
            const bool direct = N <= g_direct_factor * (L < 32 ? 32 : (L > 512 ? 512 : L));
            tomo_group_geometry(B, L, &G, &ctas);
            if (!direct && ws_bytes >= tomo_group_ws_bytes(L, G, ctas)) {
                TomoGroupArgs g;
                g.V = V; g.est = est; g.n_used = n_used; g.err = err; g.ws = (double*)ws; g.B = B; g.d = d; g.L = L;
                g.logL = logL; g.G = G; g.norm_inf = norm_inf; g.attempt = sched_len - 1; g.N = N; g.seed = seed;
                g.vec_offset = vec_offset;
                return tomo_group_launch(g, ctas, g_group_threads, stream);
            }
        }
        return sq_tomography_batched(V, B, d, delta, norm_inf, stop, sched_host, sched_len, seed, vec_offset, est,
                                     n_used, err, nullptr, nullptr, stream_);
    }
    SQ_REQUIRE(V && est && n_used && err && sched_host, "sq_tomography_batched_ws: null pointer");
    SQ_REQUIRE(B >= 0 && B <= 65535, "sq_tomography_batched_ws: B=%lld outside [0, 65535] for d > %d", B,
               sq_tomography_max_fused_dim());
    SQ_REQUIRE(sched_len >= 1 && sched_len <= 128, "sq_tomography_batched_ws: sched_len=%d outside [1,128]", sched_len);
    SQ_REQUIRE(delta > 0.0, "sq_tomography_batched_ws: delta must be > 0");
    if (B == 0) return SQ_OK;
    TomoBigArgs a;
    int rc = tomo_big_prepare(a, V, B, d, delta, norm_inf, stop, seed, vec_offset, est, n_used, err, ws, ws_bytes,
                              "sq_tomography_batched_ws");
    if (rc) return rc;
    a.early = (g_early && stop && sched_len > 1) ? 1 : 0;
    rc = tomo_big_phase(a, 0, 0, a.nblk, stream);
    if (rc) return rc;
    for (int at = 0; at < sched_len; ++at) {
        SQ_REQUIRE(sched_host[at] >= 1.0 && sched_host[at] < 9.0e15, "sq_tomography_batched_ws: schedule entry out of range");
        a.at = at; a.N = sched_host[at];
        a.may_skip = (a.early && at + 1 < sched_len) ? 1 : 0;
        for (int phase = 1; phase <= 3; ++phase) {
            rc = tomo_big_phase(a, phase, 0, a.nblk, stream);
            if (rc) return rc;
        }
    }
    return SQ_OK;
}

extern "C" int sq_tomography_workspace_layout(long long B, int d, long long out[8]) {
    SQ_REQUIRE(out && B >= 1 && d > sq_tomography_max_fused_dim(), "sq_tomography_workspace_layout: needs B >= 1 and d > %d",
               sq_tomography_max_fused_dim());
    int nblk, NB, logNB;
    size_t off[11], total;
    tomo_big_layout(B, d, &nblk, &NB, &logNB, off, &total);
    out[0] = nblk;
    out[1] = TB;
    for (int i = 0; i < 5; ++i) out[2 + i] = (long long)off[1 + i];   // bm_a, bm_b, bcnt, errp, done
    out[7] = (long long)total;
    return SQ_OK;
}

extern "C" int sq_tomography_sharded_phase(const double* V, long long B, int d, double delta, int norm_inf, int stop,
                                           double n_measure, int attempt, unsigned long long seed,
                                           unsigned long long vec_offset, double* est, double* n_used, double* err,
                                           void* ws, size_t ws_bytes, int blk_lo, int blk_hi, int phase,
                                           void* stream_) {
    cudaStream_t stream = (cudaStream_t)stream_;
    SQ_REQUIRE(V && est && n_used && err, "sq_tomography_sharded_phase: null pointer");
    SQ_REQUIRE(d > sq_tomography_max_fused_dim(), "sq_tomography_sharded_phase: d=%d fits one CTA; shard by vector instead", d);
    SQ_REQUIRE(B >= 1 && B <= 65535, "sq_tomography_sharded_phase: B=%lld outside [1, 65535]", B);
    SQ_REQUIRE(phase >= 0 && phase <= 3, "sq_tomography_sharded_phase: phase must be 0..3");
    SQ_REQUIRE(delta > 0.0 && attempt >= 0 && attempt < 128, "sq_tomography_sharded_phase: bad delta / attempt");
    SQ_REQUIRE(phase == 0 || (n_measure >= 1.0 && n_measure < 9.0e15), "sq_tomography_sharded_phase: n_measure out of range");
    TomoBigArgs a;
    int rc = tomo_big_prepare(a, V, B, d, delta, norm_inf, stop, seed, vec_offset, est, n_used, err, ws, ws_bytes,
                              "sq_tomography_sharded_phase");
    if (rc) return rc;
    SQ_REQUIRE(blk_lo >= 0 && blk_lo <= blk_hi && blk_hi <= a.nblk, "sq_tomography_sharded_phase: block range [%d, %d) outside [0, %d]",
               blk_lo, blk_hi, a.nblk);
    a.at = attempt; a.N = phase == 0 ? 1.0 : n_measure;
    return tomo_big_phase(a, phase, blk_lo, blk_hi, stream);
}

// ---- incremental mode with speculative attempt lanes -----------------------------------------
__global__ void tomo_fill_int_kernel(int* p, long long n, int v) {
    long long i = (long long)blockIdx.x * blockDim.x + threadIdx.x;
    if (i < n) p[i] = v;
}

extern "C" int sq_tomography_batched_spec(const double* V, long long B, int d, double delta, int norm_inf,
                                          const double* sched_host, int sched_len, int lanes, unsigned long long seed,
                                          unsigned long long vec_offset, double* est, double* n_used, double* err,
                                          int* attempts, int* stop_idx, void* stream_) {
    cudaStream_t stream = (cudaStream_t)stream_;
    SQ_REQUIRE(V && est && n_used && err && sched_host && stop_idx, "sq_tomography_batched_spec: null pointer");
    SQ_REQUIRE(B >= 0 && B <= 65535 && d >= 1, "sq_tomography_batched_spec: bad shape B=%lld d=%d", B, d);
    SQ_REQUIRE(sched_len >= 1 && sched_len <= 128, "sq_tomography_batched_spec: sched_len=%d outside [1,128]", sched_len);
    SQ_REQUIRE(lanes >= 1 && lanes <= 64, "sq_tomography_batched_spec: lanes=%d outside [1,64]", lanes);
    SQ_REQUIRE(delta > 0.0, "sq_tomography_batched_spec: delta must be > 0");
    if (d > sq_tomography_max_fused_dim()) {
        set_error("sq_tomography_batched_spec: d=%d exceeds the fused-path limit %d", d, sq_tomography_max_fused_dim());
        return SQ_ERR_UNSUPPORTED;
    }
    if (B == 0) return SQ_OK;
    TomoArgs a;
    a.V = V; a.est = est; a.n_used = n_used; a.err = err; a.trace = nullptr; a.attempts = attempts;
    a.stop_idx = stop_idx; a.lane_done = nullptr;
    for (int i = 0; i < sched_len; ++i) {
        SQ_REQUIRE(sched_host[i] >= 1.0 && sched_host[i] < 9.0e15, "sq_tomography_batched_spec: schedule entry out of range");
        a.sched[i] = sched_host[i];
    }
    a.B = B; a.d = d;
    int L = 2, logL = 1;
    while (L < d) { L <<= 1; ++logL; }
    a.L = L; a.logL = logL; a.sched_len = sched_len; a.norm_inf = norm_inf; a.stop = 1; a.delta = delta;
    a.seed = seed; a.vec_offset = vec_offset;
    tomo_fill_int_kernel<<<(unsigned)((B + 255) / 256), 256, 0, stream>>>(stop_idx, B, sched_len);
    a.spec = 1;
    int rc = tomo_launch(a, B, lanes, stream);
    if (rc) return rc;
    a.spec = 2;
    return tomo_launch(a, B, 1, stream);
}

// ---- the same in one launch ---------------------------------------------------------------------
__global__ void tomo_spec_state_kernel(int* state, long long B) {
    long long i = (long long)blockIdx.x * blockDim.x + threadIdx.x;
    if (i < B) { state[i] = 0x7fffffff; state[B + i] = 0; }
}

extern "C" int sq_tomography_spec_state_init(int* state, long long B, void* stream_) {
    cudaStream_t stream = (cudaStream_t)stream_;
    SQ_REQUIRE(state && B >= 1 && B <= 65535, "sq_tomography_spec_state_init: bad arguments");
    tomo_spec_state_kernel<<<(unsigned)((B + 255) / 256), 256, 0, stream>>>(state, B);
    SQ_LAUNCH_CHECK("tomo_spec_state_kernel");
    return SQ_OK;
}

extern "C" int sq_tomography_batched_spec_fused(const double* V, long long B, int d, double delta, int norm_inf,
                                                const double* sched_host, int sched_len, int lanes,
                                                unsigned long long seed, unsigned long long vec_offset, double* est,
                                                double* n_used, double* err, int* attempts, int* state, void* stream_) {
    cudaStream_t stream = (cudaStream_t)stream_;
    SQ_REQUIRE(V && est && n_used && err && sched_host && state, "sq_tomography_batched_spec_fused: null pointer");
    SQ_REQUIRE(B >= 0 && B <= 65535 && d >= 1, "sq_tomography_batched_spec_fused: bad shape B=%lld d=%d", B, d);
    SQ_REQUIRE(sched_len >= 1 && sched_len <= 128, "sq_tomography_batched_spec_fused: sched_len=%d outside [1,128]", sched_len);
    SQ_REQUIRE(lanes >= 1 && lanes <= 64, "sq_tomography_batched_spec_fused: lanes=%d outside [1,64]", lanes);
    SQ_REQUIRE(delta > 0.0, "sq_tomography_batched_spec_fused: delta must be > 0");
    if (d > sq_tomography_max_fused_dim()) {
        set_error("sq_tomography_batched_spec_fused: d=%d exceeds the fused-path limit %d", d, sq_tomography_max_fused_dim());
        return SQ_ERR_UNSUPPORTED;
    }
    if (B == 0) return SQ_OK;
    TomoArgs a;
    a.V = V; a.est = est; a.n_used = n_used; a.err = err; a.trace = nullptr; a.attempts = attempts;
    a.stop_idx = state; a.lane_done = state + B;
    for (int i = 0; i < sched_len; ++i) {
        SQ_REQUIRE(sched_host[i] >= 1.0 && sched_host[i] < 9.0e15, "sq_tomography_batched_spec_fused: schedule entry out of range");
        a.sched[i] = sched_host[i];
    }
    a.B = B; a.d = d;
    int L = 2, logL = 1;
    while (L < d) { L <<= 1; ++logL; }
    a.L = L; a.logL = logL; a.sched_len = sched_len; a.norm_inf = norm_inf; a.stop = 1; a.delta = delta;
    a.seed = seed; a.vec_offset = vec_offset;
    a.spec = 3;
    return tomo_launch(a, B, lanes, stream);
}
