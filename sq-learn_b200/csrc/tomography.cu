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

namespace sq {

struct TomoArgs {
    const double* V;      // [B, d] row-major
    double* est;          // [B, d]
    double* n_used;       // [B] measurements of the returned attempt
    double* err;          // [B] ||V - est|| of the returned attempt
    double* trace;        // optional [B, sched_len, d]: the estimate of every attempt made
    int* attempts;        // optional [B]: number of attempts made
    double sched[128];    // attempt sizes N (integers stored as doubles), by value
    long long B;
    int d;
    int L;                // leaves padded to a power of two
    int logL;
    int sched_len;
    int norm_inf;         // 0: L2, 1: L-infinity
    int stop;             // stop at the first attempt with err <= delta
    double delta;
    unsigned long long seed;
    unsigned long long vec_offset;  // global id of vector 0 (sharding / call offset)
};

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

// top-down conditional binomial splitting; cnt[1] must be set by the caller
__device__ __forceinline__ void split_tree(const double* mass, double* cnt, int logL, int tid, int T,
                                           unsigned long long seed, unsigned long long vec, unsigned stream) {
    for (int lev = 0; lev < logL; ++lev) {
        __syncthreads();
        int first = 1 << lev;
        for (int i = first + tid; i < 2 * first; i += T) {
            double n = cnt[i];
            double left = 0.0;
            if (n > 0.0) {
                Philox rng(seed, vec, stream, (unsigned)i);
                left = binomial_split(n, mass[2 * i], mass[2 * i + 1], rng);
            }
            cnt[2 * i] = left;
            cnt[2 * i + 1] = n - left;
        }
    }
    __syncthreads();
}

__global__ void tomo_fused_kernel(TomoArgs a) {
    extern __shared__ double sm[];
    const int L = a.L, d = a.d, tid = threadIdx.x, T = blockDim.x;
    double* v = sm;            // [L]
    double* P = v + L;         // [L]
    double* mass = P + L;      // [2L]
    double* cnt = mass + 2 * L;  // [2L]
    double* red = cnt + 2 * L;   // [32]

    const long long b = blockIdx.x;
    const unsigned long long vec = a.vec_offset + (unsigned long long)b;
    const double* Vg = a.V + b * (long long)d;

    double ss = 0.0;
    for (int i = tid; i < L; i += T) {
        double x = (i < d) ? Vg[i] : 0.0;
        v[i] = x;
        ss += x * x;
    }
    ss = block_sum(ss, red, tid, T);
    double nrm = sqrt(ss);
    // np.isclose(nrm, 1, rtol=1e-2): |nrm - 1| <= 1e-8 + 1e-2
    bool keep = fabs(nrm - 1.0) <= (1e-8 + 1e-2);
    if (!(nrm > 0.0)) {
        // zero (or NaN) vector: the reference divides by zero; we return zeros
        for (int i = tid; i < d; i += T) a.est[b * (long long)d + i] = 0.0;
        if (tid == 0) { a.n_used[b] = 0.0; a.err[b] = 0.0; if (a.attempts) a.attempts[b] = 0; }
        return;
    }
    if (!keep) {
        for (int i = tid; i < L; i += T) v[i] = v[i] / nrm;
    }
    __syncthreads();

    double n_last = 0.0, err_last = 0.0;
    int n_att = 0;
    for (int at = 0; at < a.sched_len; ++at) {
        const double N = a.sched[at];
        // ---- stage 1 ----
        for (int i = tid; i < L; i += T) mass[L + i] = v[i] * v[i];
        build_tree(mass, L, a.logL, tid, T);
        if (tid == 0) cnt[1] = N;
        split_tree(mass, cnt, a.logL, tid, T, a.seed, vec, (unsigned)(2 * at));
        // ---- stage 2 ----
        double sminus = 0.0;
        for (int i = tid; i < L; i += T) {
            double p = sqrt(cnt[L + i] / N);
            P[i] = p;
            double ap = 0.5 * (v[i] + p), am = 0.5 * (v[i] - p);
            mass[L + i] = ap * ap;
            sminus += am * am;
        }
        sminus = block_sum(sminus, red, tid, T);
        build_tree(mass, L, a.logL, tid, T);
        if (tid == 0) {
            Philox rng(a.seed, vec, (unsigned)(2 * at + 1), 0u);
            cnt[1] = binomial_split(N, mass[1], sminus, rng);
        }
        split_tree(mass, cnt, a.logL, tid, T, a.seed, vec, (unsigned)(2 * at + 1));
        // ---- sign rule + error ----
        double e = 0.0;
        for (int i = tid; i < L; i += T) {
            double p = P[i];
            double est = (cnt[L + i] > 0.4 * p * p * N) ? p : -p;
            P[i] = est;
            double df = v[i] - est;
            if (i < d) e = a.norm_inf ? fmax(e, fabs(df)) : e + df * df;
        }
        e = a.norm_inf ? block_max(e, red, tid, T) : sqrt(block_sum(e, red, tid, T));
        n_last = N;
        err_last = e;
        n_att = at + 1;
        if (a.trace) {
            double* tr = a.trace + (b * a.sched_len + at) * (long long)d;
            for (int i = tid; i < d; i += T) tr[i] = P[i];
        }
        if (a.stop && e <= a.delta) break;  // e is identical in every thread
    }
    __syncthreads();
    for (int i = tid; i < d; i += T) a.est[b * (long long)d + i] = P[i];
    if (tid == 0) { a.n_used[b] = n_last; a.err[b] = err_last; if (a.attempts) a.attempts[b] = n_att; }
}

}  // namespace sq

// ---------------------------------------------------------------------------
// C ABI
// ---------------------------------------------------------------------------
using namespace sq;

extern "C" int sq_tomography_max_fused_dim(void) { return 4096; }

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
    for (int i = 0; i < sched_len; ++i) {
        SQ_REQUIRE(sched_host[i] >= 1.0 && sched_host[i] < 9.0e15, "sq_tomography_batched: schedule entry out of range");
        a.sched[i] = sched_host[i];
    }
    a.B = B; a.d = d;
    int L = 2, logL = 1;
    while (L < d) { L <<= 1; ++logL; }
    a.L = L; a.logL = logL; a.sched_len = sched_len; a.norm_inf = norm_inf; a.stop = stop; a.delta = delta;
    a.seed = seed; a.vec_offset = vec_offset;
    int threads = L < 32 ? 32 : (L > 512 ? 512 : L);
    size_t smem = (size_t)(6 * L + 32) * sizeof(double);
    SQ_CUDA(cudaFuncSetAttribute(tomo_fused_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem));
    // gridDim.x limit is 2^31-1; batches beyond that are split
    long long done = 0;
    while (done < B) {
        long long nb = B - done > 1000000000LL ? 1000000000LL : B - done;
        TomoArgs c = a;
        c.V = V + done * d; c.est = est + done * d; c.n_used = n_used + done; c.err = err + done;
        if (trace) c.trace = trace + done * sched_len * d;
        if (attempts) c.attempts = attempts + done;
        c.vec_offset = vec_offset + (unsigned long long)done;
        tomo_fused_kernel<<<(unsigned)nb, threads, smem, stream>>>(c);
        SQ_LAUNCH_CHECK("tomo_fused_kernel");
        done += nb;
    }
    return SQ_OK;
}
