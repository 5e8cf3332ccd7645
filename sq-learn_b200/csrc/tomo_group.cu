// K1g — batched tomography, one attempt, many vectors: a CTA works on a GROUP of vectors at once.
//
// Same algorithm and the same draws as tomo_fused_kernel (tomography.cu; reference: real_tomography,
// sklearn/QuantumUtility/Utility.py:259-402, the single-shot branch :365-398): two multinomial stages drawn by
// conditional binomial splitting down a sum tree, the sign rule, the error.  What differs is who works on
// what.  tomo_fused_kernel gives a vector to a CTA and walks the tree level by level; a level of the upper tree
// has 1, 2, 4 ... nodes, so most lanes wait at barriers most of the time (ncu, round 1: 13 of 32 lanes active,
// 3.4 warps stalled on the barrier per issue).  Here the unit of a level is the group: with G vectors, level l
// has G * 2^l nodes, one thread per (vector, node), so a level of the upper tree already fills warps, and the
// nodes whose first candidate is rejected (binomial_head) are compacted over the whole level before the
// logarithms of the full acceptance test run.  The trees of a group (G * 2.5 L doubles) do not fit shared
// memory, so they live in a caller-provided workspace that stays L2-resident (<= 640 KB per CTA).
//
// Every floating-point addition happens in the order tomo_fused_kernel uses (pairwise tree sums; the strided
// halving of canon_sum) and every node draws from the same Philox stream (seed, vector, 2*attempt + stage,
// node), so est / n_used / err are bit-identical to tomo_fused_kernel's, whatever G and the grid.
#include "binomial.cuh"
#include "common.cuh"
#include "tomo_group.cuh"

namespace sq {

static const int GCH = 8;            // nodes per thread and chunk of a level
static const int GMAXV = 1024;       // most vectors in a group
static const int GBMAX = 32;         // most (vector, block) members in a group of the long-vector kernel

// sum_{i<L} f(v, i) for every vector v of the group, in canon_sum's order (tomography.cu): partial[i] = f(i) +
// f(i + L/2), then partial[i] += partial[i + s] for s = L/4, L/8, ..., 1.  Partials of vector v at scr + v * stride
// (L/2 doubles).  Result in out[v].
template <int GT, class F>
__device__ __forceinline__ void group_canon_sum(F f, double* scr, int stride, double* out, int Gc, int L, int logL,
                                                int tid) {
    const int h = L >> 1, logh = logL - 1;
    for (int q = tid; q < Gc * h; q += GT) {
        const int v = q >> logh, i = q & (h - 1);
        const double lo = f(v, i), hi = f(v, i + h);
        scr[v * stride + i] = lo + hi;
    }
    __syncthreads();
    // one warp per vector from here on: the strided halving needs only warp-level ordering
    const int lane = tid & 31, warp = tid >> 5;
    for (int v = warp; v < Gc; v += GT / 32) {
        double* p = scr + v * stride;
        int s = h >> 1;
        for (; s >= 32; s >>= 1) {
            for (int i = lane; i < s; i += 32) p[i] += p[i + s];
            __syncwarp();
        }
        double x = (lane < 2 * s || (s == 0 && lane == 0)) ? p[lane] : 0.0;
        for (; s >= 1; s >>= 1) x += __shfl_down_sync(0xffffffffu, x, s);
        if (lane == 0) out[v] = x;
    }
    __syncthreads();
}

// tree[v][i] = tree[v][2i] + tree[v][2i+1], bottom-up, for every vector of the group
template <int GT>
__device__ __forceinline__ void group_build(double* ws, int wsv, int Gc, int L, int logL, int tid) {
    for (int lev = logL - 1; lev >= 0; --lev) {
        __syncthreads();
        const int first = 1 << lev;
        for (int q = tid; q < Gc << lev; q += GT) {
            const int v = q >> lev, i = first + (q & (first - 1));
            double* t = ws + v * wsv;
            const double2 c = *reinterpret_cast<const double2*>(t + 2 * i);
            t[i] = c.x + c.y;
        }
    }
    __syncthreads();
}

// top-down splitting of every tree of the group; tree[v][1] holds the count to split.  Counts overwrite the
// masses in place (a node reads its two child masses right before it stores the two child counts).
// vec_of(v) / key_of(v): the Philox stream identity of member v's tree (vector id; node-id offset)
template <int GT, class VecOf, class KeyOf>
__device__ __forceinline__ void group_split(double* ws, int wsv, int Gc, int logL, int tid, unsigned long long seed,
                                            VecOf vec_of, KeyOf key_of, unsigned stream, int* list, int* ctr) {
    const int lane = tid & 31;
    int flip = 0;                                  // two pairs of counters, used by alternate chunks
    if (tid == 0) { ctr[0] = 0; ctr[1] = 0; ctr[2] = 0; ctr[3] = 0; }
    for (int lev = 0; lev < logL; ++lev) {
        const int first = 1 << lev;
        const int total = Gc << lev;
        for (int base = 0; base < total; base += GT * GCH) {
            __syncthreads();                       // previous level / chunk complete; list free
            int* ctr_cur = ctr + 2 * flip;
            if (tid == 0) { ctr[2 * (flip ^ 1)] = 0; ctr[2 * (flip ^ 1) + 1] = 0; }
            flip ^= 1;
#pragma unroll 1
            for (int c = 0; c < GCH; ++c) {
                const int q = base + c * GT + tid;
                int st = BINOM_DONE;
                if (q < total) {
                    const int v = q >> lev, i = first + (q & (first - 1));
                    double* t = ws + v * wsv;
                    const double n = t[i];
                    double left = 0.0;
                    if (n > 0.0) {
                        const double2 m = *reinterpret_cast<const double2*>(t + 2 * i);
                        Philox rng(seed, vec_of(v), stream, key_of(v) + (unsigned)i);
                        st = binomial_head(n, m.x, m.y, rng, left);
                    }
                    if (st == BINOM_DONE) *reinterpret_cast<double2*>(t + 2 * i) = make_double2(left, n - left);
                }
                const unsigned m1 = __ballot_sync(0xffffffffu, st == BINOM_NEEDS_BINV);
                const unsigned m2 = __ballot_sync(0xffffffffu, st == BINOM_NEEDS_BTRS);
                if (m1 | m2) {
                    int b1 = 0, b2 = 0;
                    if (lane == 0) {
                        if (m1) b1 = atomicAdd(&ctr_cur[0], __popc(m1));
                        if (m2) b2 = atomicAdd(&ctr_cur[1], __popc(m2));
                    }
                    b1 = __shfl_sync(0xffffffffu, b1, 0);
                    b2 = __shfl_sync(0xffffffffu, b2, 0);
                    const unsigned lt = (1u << lane) - 1u;
                    if (st == BINOM_NEEDS_BINV) list[b1 + __popc(m1 & lt)] = q - base;
                    if (st == BINOM_NEEDS_BTRS) list[GT * GCH - 1 - (b2 + __popc(m2 & lt))] = q - base;
                }
            }
            __syncthreads();
            const int nf = ctr_cur[0], nb = ctr_cur[1];
            for (int j = tid; j < nf + nb; j += GT) {
                const int q = base + (j < nf ? list[j] : list[GT * GCH - 1 - (j - nf)]);
                const int v = q >> lev, i = first + (q & (first - 1));
                double* t = ws + v * wsv;
                const double n = t[i];
                const double2 m = *reinterpret_cast<const double2*>(t + 2 * i);
                Philox rng(seed, vec_of(v), stream, key_of(v) + (unsigned)i);
                const double left = binomial_split(n, m.x, m.y, rng);
                *reinterpret_cast<double2*>(t + 2 * i) = make_double2(left, n - left);
            }
        }
    }
    __syncthreads();
}

template <int GT>
__global__ void __launch_bounds__(GT, 1024 / GT) tomo_group_kernel(TomoGroupArgs a) {
    __shared__ int list[GT * GCH];
    __shared__ int ctr[4];
    __shared__ double s_div[GMAXV];      // divisor of the vector (1 when its norm is within rtol 1e-2 of 1), 0 = zero vector
    __shared__ double s_aux[GMAXV];      // sminus of stage 2, then the error
    __shared__ int s_rescale;
    const int tid = threadIdx.x, L = a.L, logL = a.logL, d = a.d, G = a.G;
    const int wsv = (int)tomo_group_wsv(L);
    double* ws = a.ws + (size_t)blockIdx.x * G * wsv;
    const double N = a.N;
    const unsigned st1 = (unsigned)(2 * a.attempt), st2 = st1 + 1u;

    for (long long g0 = (long long)blockIdx.x * G; g0 < a.B; g0 += (long long)gridDim.x * G) {
        const int Gc = (int)((a.B - g0 < G) ? (a.B - g0) : G);
        const double* Vg = a.V + g0 * d;
        double* Eg = a.est + g0 * d;
        const unsigned long long vec0 = a.vec_offset + (unsigned long long)g0;
        __syncthreads();                  // the previous group is done with the shared arrays
        if (tid == 0) s_rescale = 0;
        // ---- ||V||^2 by the sum tree ----
        for (int q = tid; q < Gc << logL; q += GT) {
            const int v = q >> logL, i = q & (L - 1);
            const double x = (i < d) ? Vg[v * d + i] : 0.0;
            ws[v * wsv + L + i] = x * x;
        }
        group_build<GT>(ws, wsv, Gc, L, logL, tid);
        for (int v = tid; v < Gc; v += GT) {
            const double nrm = sqrt(ws[v * wsv + 1]);
            const bool keep = fabs(nrm - 1.0) <= (1e-8 + 1e-2);      // np.isclose(nrm, 1, rtol=1e-2)
            s_div[v] = !(nrm > 0.0) ? 0.0 : (keep ? 1.0 : nrm);
            if (nrm > 0.0 && !keep) s_rescale = 1;
        }
        __syncthreads();
        auto vin = [&](int v, int i) {
            const double dv = s_div[v];
            if (i >= d || dv == 0.0) return 0.0;
            const double x = Vg[v * d + i];
            return dv == 1.0 ? x : x / dv;
        };
        if (s_rescale) {                  // (uniform) some vector is normalised first: its tree is rebuilt
            for (int q = tid; q < Gc << logL; q += GT) {
                const int v = q >> logL, i = q & (L - 1);
                const double x = vin(v, i);
                ws[v * wsv + L + i] = x * x;
            }
            group_build<GT>(ws, wsv, Gc, L, logL, tid);
        }
#pragma unroll 1
        for (int stage = 0; stage < 2; ++stage) {
            if (stage == 0) {
                // ---- stage 1: N measurements over the masses x_i^2 ----
                for (int v = tid; v < Gc; v += GT) ws[v * wsv + 1] = (s_div[v] == 0.0) ? 0.0 : N;
            } else {
                // ---- stage 2: P = sqrt(n_i / N) (parked in est), masses ((x + P)/2)^2, lump sum of ((x - P)/2)^2 ----
                group_canon_sum<GT>([&](int v, int i) {
                    double* t = ws + v * wsv;
                    const double p = sqrt(t[L + i] / N), x = vin(v, i);
                    if (i < d) Eg[v * d + i] = p;
                    const double ap = 0.5 * (x + p), am = 0.5 * (x - p);
                    t[L + i] = ap * ap;
                    return am * am;
                }, ws + 2 * L, wsv, s_aux, Gc, L, logL, tid);
                group_build<GT>(ws, wsv, Gc, L, logL, tid);
                for (int v = tid; v < Gc; v += GT) {
                    double* t = ws + v * wsv;
                    double c = 0.0;
                    if (s_div[v] != 0.0) {
                        Philox rng(a.seed, vec0 + (unsigned long long)v, st2, 0u);
                        c = binomial_split(N, t[1], s_aux[v], rng);
                    }
                    t[1] = c;
                }
            }
            group_split<GT>(ws, wsv, Gc, logL, tid, a.seed, [&](int v) { return vec0 + (unsigned long long)v; },
                            [](int) { return 0u; }, stage == 0 ? st1 : st2, list, ctr);
        }
        // ---- sign rule, estimate, error ----
        auto sign_and_diff = [&](int v, int i) {
            if (i >= d) return 0.0;
            const double p = Eg[v * d + i];
            const double c2 = ws[v * wsv + L + i];
            const double est = (s_div[v] == 0.0) ? 0.0 : ((c2 > tomo_sign_threshold(p, N)) ? p : -p);
            Eg[v * d + i] = est;
            return vin(v, i) - est;
        };
        if (a.norm_inf) {
            // one warp per vector
            const int lane = tid & 31, warp = tid >> 5;
            for (int v = warp; v < Gc; v += GT / 32) {
                double m = 0.0;
                for (int i = lane; i < d; i += 32) m = fmax(m, fabs(sign_and_diff(v, i)));
                m = warp_max(m);
                if (lane == 0) s_aux[v] = m;
            }
            __syncthreads();
        } else {
            group_canon_sum<GT>([&](int v, int i) { const double df = sign_and_diff(v, i); return df * df; },
                            ws + 2 * L, wsv, s_aux, Gc, L, logL, tid);
        }
        for (int v = tid; v < Gc; v += GT) {
            const bool zero = s_div[v] == 0.0;
            a.n_used[g0 + v] = zero ? 0.0 : N;
            a.err[g0 + v] = zero ? 0.0 : (a.norm_inf ? s_aux[v] : sqrt(s_aux[v]));
        }
    }
}

// ---------------------------------------------------------------------------------------------------------
// Long vectors (d > 4096): the leaf blocks of tomography.cu's block path, handled the same way.  A member of a
// group is one (vector, block of 1024 leaves); its tree lives in the workspace, its root count comes from the top
// kernel, its node streams are keyed (vector, 4*attempt + {1, 3}, block * 2048 + node) exactly as
// tomo_big_block_kernel keys them.  Per-member sums (masses of stage 2, the sign-free lower bound, the error) are
// taken by one warp per member in an order fixed by the block alone, so nothing depends on the grouping.
template <int GT>
__global__ void __launch_bounds__(GT, 1024 / GT) tomo_blockgroup_kernel(TomoBlockGroupArgs a) {
    constexpr int TBL = 1024, TBLOG = 10;
    __shared__ int list[GT * GCH];
    __shared__ int ctr[4];
    __shared__ long long s_b[GBMAX];
    __shared__ int s_blk[GBMAX];
    __shared__ double s_div[GBMAX];      // 0: member idle (vector done / hopeless attempt / stage 2 skipped / zero vector)
    __shared__ double s_cnt[GBMAX];      // measurements that fall in the member's block
    __shared__ int s_act[GBMAX];         // members with measurements to split: slot -> member
    __shared__ int s_slot[GBMAX];        // member -> slot (-1: none; all its leaf counts are zero)
    __shared__ int s_nact;
    const int tid = threadIdx.x, lane = tid & 31, warp = tid >> 5, G = a.G;
    const int wsv = 2 * TBL;
    double* ws = a.gws + (size_t)blockIdx.x * G * wsv;
    const double N = a.N;
    const long long total = a.B * (long long)a.nown;
    const unsigned stream = (unsigned)(4 * a.at + (a.stage == 1 ? 1 : 3));
    for (long long g0 = (long long)blockIdx.x * G; g0 < total; g0 += (long long)gridDim.x * G) {
        const int Gc = (int)((total - g0 < G) ? (total - g0) : G);
        __syncthreads();
        if (tid == 0) {
            // (G <= 32 members: one thread lists them in order, so the slots do not depend on timing)
            int na = 0;
            for (int v = 0; v < Gc; ++v) {
                const long long m = g0 + v, b = m / a.nown;
                const int blk = a.blk_lo + (int)(m - b * a.nown);
                const double ss = a.sumsq[b], nrm = sqrt(ss);
                const bool zero = !(nrm > 0.0), keep = fabs(nrm - 1.0) <= (1e-8 + 1e-2);
                const bool idle = zero || a.done[b] || (a.may_skip && N < a.nhop[b]) || (a.stage == 2 && a.skip2[b]);
                s_b[v] = b; s_blk[v] = blk;
                s_div[v] = idle ? 0.0 : (keep ? 1.0 : nrm);
                const double c = idle ? 0.0 : a.bcnt[b * a.nblk + blk];
                s_cnt[v] = c;
                s_slot[v] = -1;
                if (c > 0.0) { s_slot[v] = na; s_act[na++] = v; }
            }
            s_nact = na;
        }
        __syncthreads();
        const int nact = s_nact;
        // ---- trees of the members that have measurements to split (slot s of the workspace) ----
        for (int q = tid; q < nact << TBLOG; q += GT) {
            const int sl = q >> TBLOG, i = q & (TBL - 1), v = s_act[sl];
            const double dv = s_div[v];
            double mass = 0.0;
            const int gidx = s_blk[v] * TBL + i;
            if (gidx < a.d) {
                const long long at = s_b[v] * (long long)a.d + gidx;
                const double x = a.V[at] / dv;
                if (a.stage == 1) mass = x * x;
                else { const double ap = 0.5 * (x + a.est[at]); mass = ap * ap; }
            }
            ws[sl * wsv + TBL + i] = mass;
        }
        if (nact > 0) {
            group_build<GT>(ws, wsv, nact, TBL, TBLOG, tid);
            for (int sl = tid; sl < nact; sl += GT) ws[sl * wsv + 1] = s_cnt[s_act[sl]];
            group_split<GT>(ws, wsv, nact, TBLOG, tid, a.seed,
                            [&](int sl) { return a.vec_offset + (unsigned long long)s_b[s_act[sl]]; },
                            [&](int sl) { return (unsigned)(s_blk[s_act[sl]] * (2 * TBL)); }, stream, list, ctr);
        }
        // ---- per-member epilogue, one warp per member ----
        for (int v = warp; v < Gc; v += GT / 32) {
            const double dv = s_div[v];
            if (dv == 0.0) continue;
            const long long b = s_b[v];
            const int blk = s_blk[v], sl = s_slot[v];
            const long long base = b * (long long)a.d + (long long)blk * TBL;
            const double* t = ws + (sl < 0 ? 0 : sl) * wsv + TBL;
            double s0 = 0.0, s1 = 0.0, s2 = 0.0;
            for (int i = lane; i < TBL; i += 32) {
                const int gidx = blk * TBL + i;
                const bool in = gidx < a.d;
                const double x = in ? a.V[base + i] / dv : 0.0;
                const double cnt = sl < 0 ? 0.0 : t[i];
                if (a.stage == 1) {
                    const double p = sqrt(cnt / N);
                    if (in) a.est[base + i] = p;
                    const double ap = 0.5 * (x + p), am = 0.5 * (x - p), gap = fabs(x) - p;
                    s0 += ap * ap;
                    s1 += am * am;
                    s2 = a.norm_inf ? fmax(s2, fabs(gap)) : s2 + gap * gap;
                } else if (in) {
                    const double p = a.est[base + i];
                    const double est = (cnt > tomo_sign_threshold(p, N)) ? p : -p;
                    a.est[base + i] = est;
                    const double df = x - est;
                    s0 = a.norm_inf ? fmax(s0, fabs(df)) : s0 + df * df;
                }
            }
            if (a.stage == 1) {
                s0 = warp_sum(s0); s1 = warp_sum(s1);
                s2 = a.norm_inf ? warp_max(s2) : warp_sum(s2);
                if (lane == 0) {
                    a.bm_a[b * a.nblk + blk] = s0; a.bm_b[b * a.nblk + blk] = s1;
                    if (a.may_skip) a.lbp[b * a.nblk + blk] = s2;
                }
            } else {
                s0 = a.norm_inf ? warp_max(s0) : warp_sum(s0);
                if (lane == 0) a.errp[b * a.nblk + blk] = s0;
            }
        }
    }
}

int tomo_blockgroup_launch(TomoBlockGroupArgs a, int ctas, cudaStream_t stream) {
    tomo_blockgroup_kernel<256><<<ctas, 256, 0, stream>>>(a);
    SQ_LAUNCH_CHECK("tomo_blockgroup_kernel");
    return SQ_OK;
}

size_t tomo_group_ws_bytes(int L, int G, int ctas) { return (size_t)ctas * G * tomo_group_wsv(L) * sizeof(double); }

// vectors per group: enough to fill 1024 threads from the upper levels on (G * L ~ 32768), not more than
// leaves every SM a group, at most GMAXV
int tomo_group_size(long long B, int L, int sms, int threads) {
    long long g = 32768 / L;
    if (g < 8) g = 8;
    const int per = 1024 / threads;               // CTAs per SM
    long long per_cta = (B + (long long)sms * per - 1) / ((long long)sms * per);
    if (g > per_cta) g = per_cta;
    if (g > GMAXV) g = GMAXV;
    if (g < 1) g = 1;
    return (int)g;
}

int tomo_group_launch(TomoGroupArgs a, int ctas, int threads, cudaStream_t stream) {
    // (512 and 1024 threads per CTA were measured too: 256, four CTAs per SM, is the fastest - DESIGN.md section 6)
    (void)threads;
    tomo_group_kernel<256><<<ctas, 256, 0, stream>>>(a);
    SQ_LAUNCH_CHECK("tomo_group_kernel");
    return SQ_OK;
}

}  // namespace sq
