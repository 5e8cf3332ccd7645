// Inner-product estimation for one (sample, centroid) pair, warp-cooperative.
//
// Reference: ipe -> median_evaluation(amplitude_estimation, gamma=0.1, Q=13)
// (sklearn/QuantumUtility/Utility.py:409-470, 473-508, 597-629), called for every pair by
// labels_estimation (sklearn/cluster/_dmeans.py:741-760).  The result only depends on the
// median of Q = 13 independent amplitude-estimation outcomes j_1..j_13 with
//     p(j) = sin^2(pi t) / (M^2 sin^2(pi (j - x) / M)),  x = M theta_a / pi,  t = frac(x),
// through a~ = sin^2(pi j / M), which is increasing in r = min(j, M - j).  So instead of 13
// draws and a sort the median itself is sampled, exactly:
//   * the 8 grid points around x ("window", offsets -3..4) are enumerated; T, the number of
//     the 13 draws that fall outside it, is Binomial(13, 1 - W) (one uniform);
//   * the T outside draws are sampled by rejection (envelope C / (d^2 - 1/4) per cell, C from
//     sin(pi y) >= y sin(pi/8) / (1/8) for y <= 1/8, sin(pi y) >= 2y beyond).  The lanes of the warp
//     pool their tail draws so that one pass serves all 32 pairs in flight;
//   * with b tail draws below the window's r-range and none inside it, the median of the 13
//     is the (7 - b)-th smallest of the n = 13 - T window draws, whose CDF at a window point is
//     P(Binomial(n, F) >= 7 - b), F the window CDF in r-order: one more uniform.
// Anything else (tiny grid, a tail draw inside the window's r-range, the median itself in a
// tail) replays the same tail draws and samples the window draws one by one.  The estimator comes
// in two builds: LEAN (grids of >= 32 points, window away from 0 and M/2; any other pair is
// flagged and left alone) and general; the E-step kernel runs the lean one over a tile and hands
// the flagged pairs to the general one together (one pair per row when the centroids are plain
// means: the row's own centroid sits at x ~ 1).  Every draw is keyed by (seed, row, column, purpose), so the estimate of a
// pair does not depend on which lane, warp, CTA or GPU computes it.
#pragma once
#include "common.cuh"
#include "fejer.cuh"

namespace sq {

static constexpr uint32_t IPE_TAG_MAIN = 0x1bE0u;  // (u0, V)
static constexpr uint32_t IPE_TAG_TAIL = 0x1bE4u;  // tail draws, 64 blocks per draw
static constexpr uint32_t IPE_TAG_WIN = 0x1bE5u;   // window draws of the replay path
static constexpr int IPE_Q = 13;                   // median_evaluation: Q for gamma = 0.1 (Utility.py:603-606)

// (d / (M sin(pi d / M)))^2 <= IPE_C1 for d <= M/8, <= 1/4 for d <= M/2
static constexpr double IPE_C1 = 0.10669417382416;

__device__ __forceinline__ double ipe_wrap(double j, double M) {
    if (j < 0.0) j += M;
    if (j >= M) j -= M;
    return j;
}

// One draw from p(j) restricted to the offsets outside [-3, 4]; requires M >= 10 and t > 0.
// Stream: (seed, rowkey, IPE_TAG_TAIL, col), blocks [64 q, 64 q + 64).
__device__ __noinline__ double ipe_tail_draw(double M, double t, double j0, uint64_t seed, uint64_t rowkey,
                                             uint32_t col, int q) {
    Philox rng(seed, rowkey, IPE_TAG_TAIL, col);
    rng.c0 = (uint32_t)q * 64u;
    // offsets: right m = 5 .. mhi (distance m - t), left m = -4 .. mlo (distance -m + t), mhi - mlo + 1 = M
    double half = floor((M - 1.0) * 0.5);
    double mlo = -half;
    if (t > 0.5 && M - 2.0 * floor(M * 0.5) != 0.0) mlo += 1.0;   // M odd
    const double mhi = mlo + M - 1.0;
    const double cntR = mhi - 4.0, cntL = -mlo - 3.0;      // cells on each side (>= 0)
    const double aR = 4.5 - t, aL = 3.5 + t;               // lower edge of the first cell
    const double n1R = fmin(fmax(floor(M * 0.125 - aR), 0.0), cntR);
    const double n1L = fmin(fmax(floor(M * 0.125 - aL), 0.0), cntL);
    const double iaR = 1.0 / aR, ibR = 1.0 / (aR + n1R), icR = 1.0 / (aR + cntR);
    const double iaL = 1.0 / aL, ibL = 1.0 / (aL + n1L), icL = 1.0 / (aL + cntL);
    const double eR1 = IPE_C1 * (iaR - ibR), eR2 = 0.25 * (ibR - icR);
    const double eL1 = IPE_C1 * (iaL - ibL), eL2 = 0.25 * (ibL - icL);
    const double tot = eR1 + eR2 + eL1 + eL2;
    if (!(tot > 0.0)) return j0;
    const double invM = 1.0 / M;
    for (int it = 0; it < 64; ++it) {
        double ua, ub;
        rng.uniform2(ua, ub);
        double u = ua * tot;
        // pick the sub-envelope; u is reused (rescaled) as the position inside it
        bool right;
        double ilo, ihi, a0, first, cells, C, e;
        if (u < eR1) { right = true; ilo = iaR; ihi = ibR; a0 = aR; first = 0.0; cells = n1R; C = IPE_C1; e = eR1; }
        else if ((u -= eR1) < eR2) { right = true; ilo = ibR; ihi = icR; a0 = aR; first = n1R; cells = cntR - n1R; C = 0.25; e = eR2; }
        else if ((u -= eR2) < eL1) { right = false; ilo = iaL; ihi = ibL; a0 = aL; first = 0.0; cells = n1L; C = IPE_C1; e = eL1; }
        else { u -= eL1; right = false; ilo = ibL; ihi = icL; a0 = aL; first = n1L; cells = cntL - n1L; C = 0.25; e = eL2; }
        if (!(cells >= 1.0)) continue;                       // rounding pushed u into an empty part
        double f = fmin(u / e, 1.0);
        double z = 1.0 / (ilo - f * (ilo - ihi));            // density ~ 1/z^2 on the part
        double i = floor(z - a0);
        i = fmin(fmax(i, first), first + cells - 1.0);
        const double d = a0 + i + 0.5;                       // |m - t| of the cell's grid point
        const double ms = M * sinpi(d * invM);
        if (ub * C * ms * ms <= d * d - 0.25) return ipe_wrap(j0 + rint(right ? d + t : t - d), M);
    }
    return j0;  // not reached: acceptance >= 0.38 per round
}

static __constant__ double kIpeInv[14] = {0.0, 1.0, 1.0 / 2, 1.0 / 3, 1.0 / 4, 1.0 / 5, 1.0 / 6, 1.0 / 7,
                                          1.0 / 8, 1.0 / 9, 1.0 / 10, 1.0 / 11, 1.0 / 12, 1.0 / 13};
// P(Binomial(n, F) >= k) = F^k sum_{i=0}^{n-k} C(k+i-1, i) (1-F)^i for 1 <= k <= 7, 0 <= n - k <= 6
// (k = 7 - b and n - k = 6 - c with b, c the tail draws below / above the window).  Straight-line code:
// term_i = term_{i-1} q (k-1+i)/i.
__device__ __forceinline__ double ipe_order_cdf(double F, int k, int nmk) {
    const double q = 1.0 - F, km1 = (double)(k - 1);
    double term = q * km1 + q, sum = 1.0 + ((1 <= nmk) ? term : 0.0);      // i = 1: (k-1+1)/1 = k
    term *= q * fma(km1, 1.0 / 2, 1.0); sum += (2 <= nmk) ? term : 0.0;
    term *= q * fma(km1, 1.0 / 3, 1.0); sum += (3 <= nmk) ? term : 0.0;
    term *= q * fma(km1, 1.0 / 4, 1.0); sum += (4 <= nmk) ? term : 0.0;
    term *= q * fma(km1, 1.0 / 5, 1.0); sum += (5 <= nmk) ? term : 0.0;
    term *= q * fma(km1, 1.0 / 6, 1.0); sum += (6 <= nmk) ? term : 0.0;
    const double F2 = F * F, F4 = F2 * F2;
    const double Fk = ((k & 1) ? F : 1.0) * ((k & 2) ? F2 : 1.0) * ((k & 4) ? F4 : 1.0);
    return Fk * sum;
}

// Replay path: the tail draws are regenerated from their keys, the window draws are sampled one by
// one and everything is sorted.  Also serves grids too small for a window (M <= 9).
__device__ __noinline__ double ipe_median_replay(double M, double x, int T, uint64_t seed, uint64_t rowkey,
                                                 uint32_t col) {
    FejerTable F;
    fejer_setup(F, M, x);
    double r[IPE_Q];
    int nr = 0;
    auto push = [&](double j) {
        double v = fmin(j, M - j);
        int p = nr++;
        while (p > 0 && r[p - 1] > v) { r[p] = r[p - 1]; --p; }
        r[p] = v;
    };
    Philox rng(seed, rowkey, IPE_TAG_WIN, col);
    if (F.brute) {
        for (int q = 0; q < IPE_Q; ++q) push(fejer_brute(F, rng.uniform()));
        return r[IPE_Q / 2];
    }
    for (int q = 0; q < T; ++q) push(ipe_tail_draw(M, F.t, F.j0, seed, rowkey, col, q));
    const double W = F.cum[2 * FEJER_W - 1];
    for (int q = T; q < IPE_Q; ++q) {
        double j = F.j0;
        if (!fejer_window(F, rng.uniform() * W, j)) j = fejer_wrap(F, (double)FEJER_W);  // u*W rounded up to W
        push(j);
    }
    return r[IPE_Q / 2];
}

// |x|^2 + |c|^2 - 2 ipe(x, c) for the pair held by each lane.  All 32 lanes of the warp must call
// it together; lanes without a pair pass valid = false.  eps is ipe's epsilon (delta / 2).
// LEAN = true keeps only the common case in the instruction stream (grid of >= 32 points, window away from
// 0 and M/2, no tail draw inside the window's range): any other pair sets defer and is to be given to the
// LEAN = false version later, which handles everything and regenerates the same keyed draws.
template <bool LEAN>
__device__ __forceinline__ double ipe_estimate_warp(double dot, double xn, double cn, double eps, bool valid,
                                                    uint64_t seed, uint64_t rowkey, uint32_t col, bool force_replay,
                                                    bool& defer) {
    const unsigned FULL = 0xffffffffu;
    const int lane = threadIdx.x & 31;
    const double PI = 3.14159265358979323846;
    const double S = xn + cn;
    enum { DONE = 0, FAST = 1, REPLAY = 2 };
    int status = DONE;
    double M = 16.0, x = 0.0, t = 0.5, j0 = 4.0, med = 0.0, rlo = 0.0, rhi = 0.0, W = 1.0, V = 0.5;
    double w[8];
#pragma unroll
    for (int i = 0; i < 8; ++i) w[i] = 0.0;
    int T = 0;
    bool sorted = true;
    defer = false;
    if (valid && S > 0.0) {
        double a = (S - 2.0 * dot) / (2.0 * S);
        if (a < 1e-15) a = 0.0;  // math.isclose(a, 0, abs_tol=1e-15) (Utility.py:487); negatives are rounding noise
        if (a > 1.0) a = 1.0;
        const double eps_a = eps * fmax(1.0, fabs(dot)) / S;
        M = ceil((PI / (2.0 * eps_a)) * (1.0 + sqrt(1.0 + 4.0 * eps_a)));
        x = M * (asin(sqrt(a)) / PI);
        j0 = floor(x);
        t = x - j0;
        if (t == 0.0) {
            med = fmin(j0, M - j0);  // weight 1 at distance 0 (Utility.py:500-501)
        } else if (M <= 9.0 || (LEAN && M < 32.0)) {
            status = REPLAY;
        } else {
            status = FAST;
            const double st = sinpi(t), s2 = st * st, invM = 1.0 / M;
            // p_i = (M sin(pi z_i / M))^2 at the offsets z_i = i - 3 - t
            double pw[8];
            if (LEAN || M >= 32.0) {
                // |pi z / M| < pi/8: M sin(pi z/M) = pi z P(u), u = (pi z/M)^2, Taylor to u^7 (next term < 2e-18)
                const double piM = PI * invM;
#pragma unroll
                for (int i = 0; i < 8; ++i) {
                    const double z = (double)(i - 3) - t, y = z * piM, u = y * y;
                    double P = -1.0 / 1307674368000.0;
                    P = fma(P, u, 1.0 / 6227020800.0);
                    P = fma(P, u, -1.0 / 39916800.0);
                    P = fma(P, u, 1.0 / 362880.0);
                    P = fma(P, u, -1.0 / 5040.0);
                    P = fma(P, u, 1.0 / 120.0);
                    P = fma(P, u, -1.0 / 6.0);
                    P = fma(P, u, 1.0);
                    const double ms = PI * z * P;
                    pw[i] = ms * ms;
                }
            } else {
#pragma unroll
                for (int i = 0; i < 8; ++i) {
                    const double ms = M * sinpi(((double)(i - 3) - t) * invM);
                    pw[i] = ms * ms;
                }
            }
            // w_i = s2 / p_i with one division: prefix and suffix products
            {
                const double a01 = pw[0] * pw[1], a23 = pw[2] * pw[3], a45 = pw[4] * pw[5], a67 = pw[6] * pw[7];
                const double lo4 = a01 * a23, hi4 = a45 * a67;
                const double r = s2 / (lo4 * hi4);
                const double rl = r * hi4, rh = r * lo4;          // s2 / lo4, s2 / hi4
                const double r01 = rl * a23, r23 = rl * a01, r45 = rh * a67, r67 = rh * a45;   // s2 / a01, ...
                w[0] = r01 * pw[1]; w[1] = r01 * pw[0]; w[2] = r23 * pw[3]; w[3] = r23 * pw[2];
                w[4] = r45 * pw[5]; w[5] = r45 * pw[4]; w[6] = r67 * pw[7]; w[7] = r67 * pw[6];
            }
            double acc = ((w[0] + w[1]) + (w[2] + w[3])) + ((w[4] + w[5]) + (w[6] + w[7]));
            W = acc;
            // r-order of the window: increasing with the offset unless it straddles 0 or M/2
            sorted = (j0 >= 3.0) && (2.0 * (j0 + 4.0) <= M);
            if (sorted) { rlo = j0 - 3.0; rhi = j0 + 4.0; }
            else if (LEAN) { status = REPLAY; }
            else {
                rlo = M; rhi = 0.0;
#pragma unroll
                for (int i = 0; i < 8; ++i) {
                    double j = ipe_wrap(j0 + (double)(i - 3), M);
                    double r = fmin(j, M - j);
                    rlo = fmin(rlo, r); rhi = fmax(rhi, r);
                }
            }
            Philox rng(seed, rowkey, IPE_TAG_MAIN, col);
            double u0;
            rng.uniform2(u0, V);
            // T ~ Binomial(13, tau), tau = 1 - W, by inversion from T = 0
            const double tau = fmax(1.0 - W, 0.0), Wc = fmin(W, 1.0);
            double W2 = Wc * Wc, W4 = W2 * W2, W8 = W4 * W4;
            double pt = W8 * W4 * Wc, cum = pt;
            const double ratio = tau / Wc;
            while (u0 >= cum && T < IPE_Q) {
                pt *= ratio * ((double)(IPE_Q - T) * kIpeInv[T + 1]);
                cum += pt;
                ++T;
            }
        }
    }

    // ---- pooled tail draws: draw idx of the warp goes to lane idx % 32 ----
    int b = 0, c = 0;
    bool mixed = false;
    {
        const int Tt = (status == FAST) ? T : 0;
        int incl = Tt;
#pragma unroll
        for (int o = 1; o < 32; o <<= 1) {
            int y = __shfl_up_sync(FULL, incl, o);
            if (lane >= o) incl += y;
        }
        const int excl = incl - Tt;
        const int total = __shfl_sync(FULL, incl, 31);
        for (int base = 0; base < total; base += 32) {
            const int idx = base + lane;
            int owner = 0;  // number of lanes whose inclusive count is <= idx
#pragma unroll
            for (int step = 16; step >= 1; step >>= 1) {
                int v = __shfl_sync(FULL, incl, (owner + step - 1) & 31);
                if (v <= idx) owner += step;
            }
            owner &= 31;
            const double oM = __shfl_sync(FULL, M, owner), ot = __shfl_sync(FULL, t, owner);
            const double oj0 = __shfl_sync(FULL, j0, owner);
            const double olo = __shfl_sync(FULL, rlo, owner), ohi = __shfl_sync(FULL, rhi, owner);
            const unsigned long long okey = __shfl_sync(FULL, (unsigned long long)rowkey, owner);
            const unsigned ocol = __shfl_sync(FULL, col, owner);
            const int q = idx - __shfl_sync(FULL, excl, owner);
            int cls = 0;
            if (idx < total) {
                double j = ipe_tail_draw(oM, ot, oj0, seed, okey, ocol, q);
                double r = fmin(j, oM - j);
                cls = r < olo ? 1 : (r > ohi ? 2 : 3);
            }
            const unsigned mb = __ballot_sync(FULL, cls == 1), mc = __ballot_sync(FULL, cls == 2),
                           mm = __ballot_sync(FULL, cls == 3);
            int s = excl - base, e = s + Tt;
            s = max(s, 0); e = min(e, 32);
            if (e > s) {
                unsigned rm = ((e - s) == 32) ? FULL : (((1u << (e - s)) - 1u) << s);
                b += __popc(mb & rm); c += __popc(mc & rm); mixed = mixed || ((mm & rm) != 0u);
            }
        }
    }

    // ---- the (7 - b)-th smallest of the n = 13 - T window draws ----
    const int k = 7 - b, n = IPE_Q - T;
    if (status == FAST && (mixed || k < 1 || k > n || force_replay)) status = REPLAY;
    {
        const bool fast = (status == FAST);
        const int kk = fast ? k : 1, nmk = fast ? n - k : 0;
        const double iW = 1.0 / W;
        const bool srt = LEAN || __all_sync(FULL, sorted || !fast);
        if (srt) {
            // cumulative window CDF in offset order; smallest i with V <= G(F_i), by bisection (F_7 = 1)
            const double F0 = w[0] * iW, F1 = F0 + w[1] * iW, F2 = F1 + w[2] * iW, F3 = F2 + w[3] * iW,
                         F4 = F3 + w[4] * iW, F5 = F4 + w[5] * iW, F6 = F5 + w[6] * iW;
            const bool h1 = V <= ipe_order_cdf(F3, kk, nmk);                       // answer in 0..3
            const bool h2 = V <= ipe_order_cdf(h1 ? F1 : F5, kk, nmk);             // lower pair of the half
            const double Fq = h1 ? (h2 ? F0 : F2) : (h2 ? F4 : F6);
            const bool h3 = V <= ipe_order_cdf(Fq, kk, nmk);
            const int i = (h1 ? 0 : 4) + (h2 ? 0 : 2) + (h3 ? 0 : 1);
            if (fast) med = rlo + (double)i;
        } else if (!LEAN) {
            double best = INFINITY;
#pragma unroll
            for (int i = 0; i < 8; ++i) {
                const double ji = ipe_wrap(j0 + (double)(i - 3), M), ri = fmin(ji, M - ji);
                double Fi = 0.0;
#pragma unroll
                for (int i2 = 0; i2 < 8; ++i2) {
                    const double j2 = ipe_wrap(j0 + (double)(i2 - 3), M), r2 = fmin(j2, M - j2);
                    Fi += (r2 <= ri) ? w[i2] : 0.0;
                }
                const double G = (Fi >= W) ? 1.0 : ipe_order_cdf(Fi * iW, kk, nmk);
                if (V <= G) best = fmin(best, ri);
            }
            if (fast) med = (best <= rhi) ? best : rhi;   // rounding in F_i at the last point
        }
    }
    if (status == REPLAY) {
        if (LEAN) { defer = true; return 0.0; }
        med = ipe_median_replay(M, x, T, seed, rowkey, col);
    }
    if (!(valid && S > 0.0)) return 0.0;
    const double sn = sinpi(med / M);
    return __dmul_rn(__dmul_rn(2.0 * S, sn), sn);   // (never fused with the caller's subtraction of |x|^2)
}

}  // namespace sq
